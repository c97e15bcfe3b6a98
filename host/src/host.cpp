// Host mirror of the reference's planner-side types above the C-ABI (see nos/host.hpp).  No planning logic lives
// here: parsing, interning, packing, one nosplan_plan call, unpacking.
#include "nos/host.hpp"

#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstring>
#include <set>
#include <unordered_map>

namespace nos {
namespace {

const char* kStatusPrefix = "nos.nebuly.com/status-gpu";          // v1alpha1.AnnotationGpuStatusPrefix
const char* kSpecPrefix = "nos.nebuly.com/spec-gpu";              // v1alpha1.AnnotationGpuSpecPrefix
const char* kLabelPartitioning = "nos.nebuly.com/gpu-partitioning";
const char* kLabelProduct = "nvidia.com/gpu.product";
const char* kLabelCount = "nvidia.com/gpu.count";
const char* kLabelMemory = "nvidia.com/gpu.memory";
const std::string kMigPrefix = "nvidia.com/mig-";
const std::string kMpsPrefix = "nvidia.com/gpu-";

bool starts_with(const std::string& s, const std::string& p) { return s.size() >= p.size() && s.compare(0, p.size(), p) == 0; }
bool is_digit(char c) { return c >= '0' && c <= '9'; }

// strconv.Atoi: optional sign, decimal digits only, must fit an int (64-bit on the reference's platforms; the path
// only ever stores small counts, so values beyond int32 are rejected here as out of range).
std::optional<int> atoi_go(const std::string& s) {
    size_t i = 0;
    if (i < s.size() && (s[i] == '+' || s[i] == '-')) i++;
    if (i == s.size()) return std::nullopt;
    long long v = 0;
    for (size_t k = i; k < s.size(); k++) {
        if (!is_digit(s[k])) return std::nullopt;
        v = v * 10 + (s[k] - '0');
        if (v > (long long)INT_MAX + 1) return std::nullopt;
    }
    if (s[0] == '-') v = -v;
    if (v > INT_MAX || v < INT_MIN) return std::nullopt;
    return (int)v;
}

std::vector<std::string> split(const std::string& s, char sep) {
    std::vector<std::string> out;
    size_t b = 0;
    for (;;) {
        size_t e = s.find(sep, b);
        out.push_back(s.substr(b, e == std::string::npos ? e : e - b));
        if (e == std::string::npos) break;
        b = e + 1;
    }
    return out;
}

std::string lower(std::string s) {
    for (auto& c : s) if (c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');
    return s;
}

// length of the digit run starting at i
size_t digits_at(const std::string& s, size_t i) { size_t j = i; while (j < s.size() && is_digit(s[j])) j++; return j - i; }

// regexp `nvidia\.com\/mig-\d+g\.\d+gb` (pkg/constant/constants.go:50), unanchored MatchString
bool matches_mig_resource(const std::string& r) {
    for (size_t at = r.find(kMigPrefix); at != std::string::npos; at = r.find(kMigPrefix, at + 1)) {
        size_t i = at + kMigPrefix.size(), d = digits_at(r, i);
        if (!d) continue;
        i += d;
        if (r.compare(i, 2, "g.") != 0) continue;
        i += 2;
        d = digits_at(r, i);
        if (!d) continue;
        i += d;
        if (r.compare(i, 2, "gb") == 0) return true;
    }
    return false;
}
// regexp `\d+g\.\d+gb` (pkg/constant/constants.go:51), unanchored: mig.ProfileName.isValid (pkg/gpu/mig/profile.go:57-59)
bool matches_mig_profile(const std::string& p) {
    for (size_t i = 0; i < p.size(); ++i) {
        size_t d = digits_at(p, i);
        if (!d) continue;
        size_t j = i + d;
        if (p.compare(j, 2, "g.") == 0) {
            size_t d2 = digits_at(p, j + 2);
            if (d2 && p.compare(j + 2 + d2, 2, "gb") == 0) return true;
        }
        i = j - 1;                       // a shorter suffix of the same digit run ends at the same place: same verdict
    }
    return false;
}
// regexp `nvidia\.com/gpu-\d+gb` (pkg/gpu/slicing/profile.go:30-31)
bool matches_mps_resource(const std::string& r) {
    for (size_t at = r.find(kMpsPrefix); at != std::string::npos; at = r.find(kMpsPrefix, at + 1)) {
        size_t i = at + kMpsPrefix.size(), d = digits_at(r, i);
        if (d && r.compare(i + d, 2, "gb") == 0) return true;
    }
    return false;
}

// mig.ExtractProfileName (pkg/gpu/mig/util.go:46-52) / slicing.ExtractProfileName (pkg/gpu/slicing/util.go:34-40)
std::optional<std::string> slice_profile_of(core::Kind kind, const std::string& r) {
    if (kind == core::Kind::MIG) {
        if (!matches_mig_resource(r)) return std::nullopt;
        return starts_with(r, kMigPrefix) ? r.substr(kMigPrefix.size()) : r;
    }
    if (!matches_mps_resource(r)) return std::nullopt;
    return starts_with(r, kMpsPrefix) ? r.substr(kMpsPrefix.size()) : r;
}

// leftmost match of `\d+<suffix>`: the first maximal digit run that is followed by the suffix
int digits_before(const std::string& s, const char* suffix) {
    size_t n = std::strlen(suffix);
    for (size_t i = 0; i < s.size();) {
        size_t d = digits_at(s, i);
        if (!d) { i++; continue; }
        if (s.compare(i + d, n, suffix) == 0) {
            long long v = 0;
            for (size_t k = i; k < i + d; k++) v = std::min<long long>(v * 10 + (s[k] - '0'), INT_MAX);
            return (int)v;
        }
        i += d;
    }
    return 0;
}
int mig_mem_gb(const std::string& p) { return digits_before(p, "gb"); }   // mig/profile.go:70-75
int mig_gi(const std::string& p) { return digits_before(p, "g"); }        // mig/profile.go:77-82
int mps_size_gb(const std::string& p) {                                     // slicing/profile.go:52-59
    std::string t = starts_with(p, kMpsPrefix) ? p.substr(kMpsPrefix.size()) : p;
    if (t.size() >= 2 && t.compare(t.size() - 2, 2, "gb") == 0) t.resize(t.size() - 2);
    return atoi_go(t).value_or(0);
}

mig::KnownGeometries g_known;
bool g_known_set = false;

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
namespace resource {
ResourceList ComputePodRequest(const v1::Pod& pod) {
    ResourceList cont, init;
    for (auto& c : pod.containers) for (auto& [k, v] : c.requests) cont[k] += v;
    for (auto& c : pod.initContainers) {
        std::set<std::string> keys;
        for (auto& [k, v] : init) keys.insert(k);
        for (auto& [k, v] : c.requests) keys.insert(k);
        for (auto& k : keys) {
            auto a = init.find(k); auto b = c.requests.find(k);
            init[k] = std::max<int64_t>(a == init.end() ? 0 : a->second, b == c.requests.end() ? 0 : b->second);
        }
    }
    ResourceList out;
    for (auto& [k, v] : cont) { auto i = init.find(k); out[k] = std::max<int64_t>(v, i == init.end() ? 0 : i->second); }
    for (auto& [k, v] : init) if (!cont.count(k)) out[k] = std::max<int64_t>(0, v);
    return out;
}
}  // namespace resource

namespace gpu {
std::optional<SpecAnnotation> ParseSpecAnnotation(const std::string& key, const std::string& value) {
    if (!starts_with(key, kSpecPrefix)) return std::nullopt;
    auto parts = split(key, '-');
    if (parts.size() != 4) return std::nullopt;
    auto q = atoi_go(value), idx = atoi_go(parts[2]);
    if (!q || !idx) return std::nullopt;
    return SpecAnnotation{parts[3], *idx, *q};
}
std::optional<StatusAnnotation> ParseStatusAnnotation(const std::string& key, const std::string& value) {
    if (!starts_with(key, kStatusPrefix)) return std::nullopt;
    auto parts = split(key, '-');
    if (parts.size() != 5) return std::nullopt;
    auto q = atoi_go(value), idx = atoi_go(parts[2]);
    if (!q || !idx) return std::nullopt;
    std::string st = lower(parts[4]);                                       // resource.ParseStatus (pkg/resource/device.go:28-39)
    Status s;
    if (st == "free") s = Status::Free; else if (st == "used") s = Status::Used; else if (st == "unknown") s = Status::Unknown;
    else return std::nullopt;
    return StatusAnnotation{parts[3], *idx, s, *q};
}
void ParseNodeAnnotations(const v1::Node& node, std::vector<StatusAnnotation>& status, std::vector<SpecAnnotation>& spec) {
    for (auto& [k, v] : node.annotations) {
        if (auto a = ParseSpecAnnotation(k, v)) { spec.push_back(*a); continue; }
        if (auto a = ParseStatusAnnotation(k, v)) { status.push_back(*a); continue; }
    }
}
std::string GetModel(const v1::Node& node) {
    auto it = node.labels.find(kLabelProduct);
    if (it == node.labels.end())
        throw Error("cannot get GPU model from node " + node.name + " labels: missing label " + kLabelProduct);
    return it->second;
}
int GetCount(const v1::Node& node) {
    auto it = node.labels.find(kLabelCount);
    if (it == node.labels.end()) throw Error(std::string("cannot get GPU count from node labels, missing label ") + kLabelCount);
    auto v = atoi_go(it->second);
    if (!v) throw Error("strconv.Atoi: parsing \"" + it->second + "\": invalid syntax");
    return *v;
}
int GetMemoryGB(const v1::Node& node) {
    auto it = node.labels.find(kLabelMemory);
    if (it == node.labels.end()) throw Error(std::string("cannot get GPU Memory GB from node labels, missing label ") + kLabelMemory);
    auto v = atoi_go(it->second);
    if (!v) throw Error("strconv.Atoi: parsing \"" + it->second + "\": invalid syntax");
    return (int)std::ceil((double)*v / 1000.0);
}
}  // namespace gpu

gpu::Geometry GPU::GetGeometry() const {
    gpu::Geometry g;
    for (auto& [p, q] : used) g[p] += q;
    for (auto& [p, q] : free) g[p] += q;
    return g;
}

namespace {
// shared body of mig.extractGPUs (node.go:65-100) and slicing.extractGPUs (node.go:48-105).  The reference iterates a
// Go map of annotated GPU indices; the canonical order is ascending index.
std::vector<GPU> extract_gpus(const v1::Node& node, const std::string& model, int count, int memoryGB) {
    std::vector<gpu::StatusAnnotation> st; std::vector<gpu::SpecAnnotation> sp;
    gpu::ParseNodeAnnotations(node, st, sp);
    std::map<int, GPU> by;
    for (auto& a : st) {
        GPU& g = by[a.Index];
        g.model = model; g.index = a.Index; g.memoryGB = memoryGB;
        if (a.status == gpu::Status::Used) g.used[a.ProfileName] = a.Quantity;
        if (a.status == gpu::Status::Free) g.free[a.ProfileName] = a.Quantity;
    }
    std::vector<GPU> out;
    for (auto& [i, g] : by) out.push_back(g);
    for (int i = (int)out.size(); i < count; i++) { GPU g; g.model = model; g.index = i; g.memoryGB = memoryGB; out.push_back(g); }
    return out;
}
}  // namespace

namespace mig {
const KnownGeometries& DefaultKnownGeometries() {
    static const KnownGeometries t = {
        {"NVIDIA-A30", {{{"4g.24gb", 1}}, {{"2g.12gb", 2}}, {{"2g.12gb", 1}, {"1g.6gb", 2}}, {{"1g.6gb", 4}}}},
        {"NVIDIA-A100-40GB-SXM4",
         {{{"7g.40gb", 1}},
          {{"4g.20gb", 1}, {"2g.10gb", 1}, {"1g.5gb", 1}},
          {{"4g.20gb", 1}, {"1g.5gb", 3}},
          {{"3g.20gb", 2}},
          {{"3g.20gb", 1}, {"2g.10gb", 1}, {"1g.5gb", 1}},
          {{"3g.20gb", 1}, {"1g.5gb", 3}},
          {{"2g.10gb", 2}, {"3g.20gb", 1}},
          {{"2g.10gb", 1}, {"1g.5gb", 2}, {"3g.20gb", 1}},
          {{"2g.10gb", 3}, {"1g.5gb", 1}},
          {{"2g.10gb", 2}, {"1g.5gb", 3}},
          {{"2g.10gb", 1}, {"1g.5gb", 5}},
          {{"1g.5gb", 7}}}},
        {"NVIDIA-A100-80GB-PCIe",
         {{{"7g.79gb", 1}},
          {{"4g.40gb", 1}, {"2g.20gb", 1}, {"1g.10gb", 1}},
          {{"4g.40gb", 1}, {"1g.10gb", 3}},
          {{"3g.40gb", 2}},
          {{"3g.40gb", 1}, {"2g.20gb", 1}, {"1g.10gb", 1}},
          {{"3g.40gb", 1}, {"1g.10gb", 3}},
          {{"2g.20gb", 2}, {"3g.20gb", 1}},
          {{"2g.10gb", 1}, {"1g.10gb", 2}, {"3g.40gb", 1}},
          {{"2g.20gb", 3}, {"1g.10gb", 1}},
          {{"2g.20gb", 2}, {"1g.10gb", 3}},
          {{"2g.20gb", 1}, {"1g.10gb", 5}},
          {{"1g.10gb", 7}}}},
    };
    return t;
}
void SetKnownGeometries(const KnownGeometries& g) {
    if (g.empty()) throw Error("no known configs provided");                 // ValidateConfigs (known_configs.go:159-181)
    for (auto& [model, list] : g)
        for (auto& geom : list)
            for (auto& [p, q] : geom) {
                if (!matches_mig_profile(p)) throw Error("invalid profile " + p);
                if (q < 1) throw Error("invalid quantity " + std::to_string(q) + " for profile " + p);
            }
    g_known = g; g_known_set = true;
}
const KnownGeometries& GetKnownGeometries() { return g_known_set ? g_known : DefaultKnownGeometries(); }

Node NewNode(const framework::NodeInfo& n) {
    if (!n.node) throw Error("node is nil");
    std::string model = gpu::GetModel(*n.node);
    int count = gpu::GetCount(*n.node);
    if (!GetKnownGeometries().count(model))                                  // mig.NewGPU (gpu.go:43-55)
        throw Error("model " + model + " is not associated with any known GPU");
    return Node{n.node->name, extract_gpus(*n.node, model, count, 0), n};
}
}  // namespace mig

namespace slicing {
Node NewNode(const framework::NodeInfo& n) {
    if (!n.node) throw Error("node is nil");
    std::string model = gpu::GetModel(*n.node);
    int count = gpu::GetCount(*n.node);
    int mem = gpu::GetMemoryGB(*n.node);
    return Node{n.node->name, extract_gpus(*n.node, model, count, mem), n};
}
}  // namespace slicing

namespace state {
bool NodePartitioning::Equal(const NodePartitioning& other) const {         // util.UnorderedEqual (pkg/util/util.go:170-198)
    if (GPUs.size() != other.GPUs.size()) return false;
    std::vector<char> visited(other.GPUs.size(), 0);
    for (auto& g : GPUs) {
        bool found = false;
        for (size_t j = 0; j < other.GPUs.size(); j++)
            if (!visited[j] && g == other.GPUs[j]) { visited[j] = 1; found = true; break; }
        if (!found) return false;
    }
    return true;
}
bool PartitioningState::Equal(const PartitioningState& other) const {
    if (size() != other.size()) return false;
    for (auto& [name, np] : *this) {
        auto it = other.find(name);
        if (!np.Equal(it == other.end() ? NodePartitioning{} : it->second)) return false;   // other[node] zero value
    }
    return true;
}
}  // namespace state

// ---------------------------------------------------------------------------------------------------------------
namespace mig {
std::vector<gpu::SpecAnnotation> GetGPUSpecAnnotationList(const state::NodePartitioning& p) {
    std::vector<gpu::SpecAnnotation> res;
    for (auto& g : p.GPUs)
        for (auto& [r, q] : g.Resources) {
            auto profile = slice_profile_of(core::Kind::MIG, r);
            if (!profile) throw Error("invalid input string, required format is nvidia\\.com\\/mig-\\d+g\\.\\d+gb");
            res.push_back(gpu::SpecAnnotation{*profile, g.GPUIndex, q});
        }
    return res;
}
std::string SpecAnnotationKey(const gpu::SpecAnnotation& a) {
    return std::string(kSpecPrefix) + "-" + std::to_string(a.Index) + "-" + a.ProfileName;
}
}  // namespace mig

namespace mps {
PluginConfig ToPluginConfig(const state::NodePartitioning& p) {
    PluginConfig cfg;
    for (auto& g : p.GPUs)
        for (auto& [r, q] : g.Resources) {
            auto profile = slice_profile_of(core::Kind::MPS, r);
            if (!profile) throw Error("invalid input string, required format is nvidia\\.com/gpu-\\d+gb");
            MPSResource m;
            m.Name = "nvidia.com/gpu";
            m.Rename = starts_with(r, "nvidia.com/") ? r.substr(11) : r;       // strings.TrimPrefix(r, constant.NvidiaResourcePrefix)
            m.MemoryGB = mps_size_gb(*profile);
            m.Devices = {std::to_string(g.GPUIndex)};
            m.Replicas = q;
            cfg.Resources.push_back(std::move(m));
        }
    return cfg;
}
}  // namespace mps

namespace capacityscheduling {
namespace {
// mig.ExtractMemoryGBFromMigFormat (pkg/gpu/mig/util.go:68-86): the resource must match the MIG pattern and hold exactly one
// `\\d+gb`; 0 where the reference returns an error (pkg/gpu/util/resource.go:70-86 skips the resource then)
int64_t mig_memory_gb_of_resource(const std::string& r) {
    if (!matches_mig_resource(r)) return 0;
    int n = 0; long long val = 0;
    for (size_t i = 0; i < r.size();) {
        const size_t d = digits_at(r, i);
        if (!d) { ++i; continue; }
        if (r.compare(i + d, 2, "gb") == 0) {
            ++n;
            val = 0;
            for (size_t k = i; k < i + d; ++k) val = std::min<long long>(val * 10 + (r[k] - '0'), INT_MAX);
            i += d + 2;
        } else i += d;
    }
    return n == 1 ? val : 0;
}
}  // namespace

nosplan_quota_pods PackedQuotaPods::c() const {
    nosplan_quota_pods p{};
    p.n_pods = (int32_t)ns.size(); p.ns = ns.data(); p.request = request.data(); p.request_mask = requestMask.data();
    p.priority = priority.data(); p.creation_ts = creationTs.data(); p.name_rank = nameRank.data();
    return p;
}
nosplan_quotas PackedQuotas::c() const {
    nosplan_quotas q{};
    q.n_quotas = (int32_t)names.size(); q.n_res = R(); q.n_namespaces = (int32_t)namespaces.size(); q.gpu_memory_col = gpuMemoryCol;
    q.ns_quota = nsQuota.data(); q.min = min.data(); q.max = max.data(); q.used_init = usedInit.data();
    q.min_mask = minMask.data(); q.max_mask = maxMask.data(); q.used_mask = usedMask.data(); q.max_enforced = maxEnforced.data();
    q.gpu_memory_coeff = gpuMemoryCoeff.data();
    return q;
}

PackedQuotas PackQuotas(const std::vector<ElasticQuotaInfo>& quotas, const std::vector<std::vector<QuotaPod>>& podSets, int nvidiaGpuMemoryGB) {
    static const char* kGpuMemory = "nos.nebuly.com/gpu-memory";
    static const char* kNvidiaGpu = "nvidia.com/gpu";
    PackedQuotas pq;
    auto fixed = [](const std::string& r) { return r == "cpu" ? 0 : r == "memory" ? 1 : r == "ephemeral-storage" ? 2 : r == "pods" ? 3 : -1; };
    std::vector<std::string> scalars;
    auto note = [&](const ResourceList& rl) {
        for (auto& [r, v] : rl) { (void)v; if (fixed(r) < 0 && std::find(scalars.begin(), scalars.end(), r) == scalars.end()) scalars.push_back(r); }
    };
    for (auto& e : quotas) { if (e.Min) note(*e.Min); if (e.Max) note(*e.Max); if (e.Used) note(*e.Used); }
    for (auto& pods : podSets) for (auto& p : pods) note(p.request);
    if (std::find(scalars.begin(), scalars.end(), kGpuMemory) == scalars.end()) scalars.push_back(kGpuMemory);
    std::sort(scalars.begin(), scalars.end());
    pq.columns = {"cpu", "memory", "ephemeral-storage", "pods"};
    pq.columns.insert(pq.columns.end(), scalars.begin(), scalars.end());
    const int R = pq.R();
    if (R > NOSPLAN_QUOTA_MAX_RES) throw Error(std::to_string(R) + " quota resource columns > NOSPLAN_QUOTA_MAX_RES");
    std::map<std::string, int> col;
    for (int c = 0; c < R; ++c) col[pq.columns[c]] = c;
    pq.gpuMemoryCol = col[kGpuMemory];
    pq.gpuMemoryCoeff.assign(R, 0);
    for (auto& [r, c] : col) {
        if (r == kNvidiaGpu) pq.gpuMemoryCoeff[c] = nvidiaGpuMemoryGB;
        else if (matches_mig_resource(r)) pq.gpuMemoryCoeff[c] = mig_memory_gb_of_resource(r);
    }
    std::map<std::string, int> nsid;
    auto intern = [&](const std::string& ns) {
        auto it = nsid.find(ns);
        if (it != nsid.end()) return it->second;
        const int id = (int)pq.namespaces.size();
        pq.namespaces.push_back(ns); nsid[ns] = id;
        return id;
    };
    for (auto& e : quotas) for (auto& ns : e.Namespaces) intern(ns);
    for (auto& pods : podSets) for (auto& p : pods) intern(p.ns);
    const size_t E = quotas.size();
    pq.nsQuota.assign(pq.namespaces.size(), -1);
    pq.min.assign(E * R, 0); pq.max.assign(E * R, 0); pq.usedInit.assign(E * R, 0);
    pq.minMask.assign(E, 0); pq.maxMask.assign(E, 0); pq.usedMask.assign(E, 0); pq.maxEnforced.assign(E, 0);
    auto fill = [&](int64_t* row, const std::optional<ResourceList>& rl) -> uint32_t {
        if (!rl) return 0;
        uint32_t m = 0;
        for (auto& [r, v] : *rl) { row[col[r]] += v; m |= 1u << col[r]; }
        return m | NOSPLAN_QUOTA_NONNIL;
    };
    for (size_t i = 0; i < E; ++i) {
        const ElasticQuotaInfo& e = quotas[i];
        pq.names.push_back(e.ResourceName);
        for (auto& ns : e.Namespaces) pq.nsQuota[nsid[ns]] = (int32_t)i;
        pq.minMask[i] = fill(&pq.min[i * R], e.Min);
        pq.maxMask[i] = fill(&pq.max[i * R], e.Max);
        pq.usedMask[i] = fill(&pq.usedInit[i * R], e.Used);
        pq.maxEnforced[i] = e.MaxEnforced ? 1 : 0;
    }
    for (auto& pods : podSets) {
        PackedQuotaPods pp;
        const size_t P = pods.size();
        pp.request.assign(P * R, 0); pp.requestMask.assign(P, 0); pp.nameRank.assign(P, 0);
        std::vector<size_t> order(P);
        for (size_t i = 0; i < P; ++i) {
            const QuotaPod& p = pods[i];
            order[i] = i;
            pp.keys.push_back(p.ns + "/" + p.name);
            pp.ns.push_back(nsid[p.ns]); pp.priority.push_back(p.priority); pp.creationTs.push_back(p.creationTimestamp);
            for (auto& [r, v] : p.request) {
                if (r == kGpuMemory) continue;
                const int c = col[r];
                pp.request[i * R + c] += v;
                if (c >= 4) pp.requestMask[i] |= 1u << c;
            }
        }
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return pods[a].name < pods[b].name; });   // byte order
        int rank = -1;
        for (size_t k = 0; k < P; ++k) {
            if (k == 0 || pods[order[k]].name != pods[order[k - 1]].name) ++rank;
            pp.nameRank[order[k]] = rank;
        }
        pq.podSets.push_back(std::move(pp));
    }
    return pq;
}
nosplan_preempt_query PackedPreemptQuery::c() const {
    nosplan_preempt_query q{};
    q.v.preemptor_ns = preemptorNs; q.v.preemptor_priority = preemptorPriority; q.v.nominated_in_eq = nominatedInEQ.data();
    q.v.n_nodes = (int32_t)nodePtr.size() - 1; q.v.node_ptr = nodePtr.data(); q.v.n_pods = (int32_t)ns.size();
    q.v.ns = ns.data(); q.v.priority = priority.data(); q.v.over_quota = overQuota.data(); q.v.request = request.data();
    q.v.request_mask = requestMask.data();
    q.pod_req = podReq.data(); q.nominated = nominated.data(); q.fit_request = fitRequest.data(); q.fit_mask = fitMask;
    q.n_pdbs = (int32_t)pdbAllowed.size(); q.node_alloc = nodeAlloc.data(); q.node_requested = nodeRequested.data();
    q.pdb_allowed = pdbAllowed.data(); q.pod_pdb_ptr = podPdbPtr.data(); q.pod_pdb = podPdb.data();
    return q;
}

PackedPreemptQuery PackPreemptQuery(const PackedQuotas& pq, const QuotaPod& preemptor, const ResourceList& nominatedInEQ,
                                    const ResourceList& nominated, const std::vector<PreemptNode>& nodes,
                                    const std::vector<int32_t>& pdbAllowed) {
    static const std::string kGpuMemory = "nos.nebuly.com/gpu-memory";
    const int R = pq.R();
    std::map<std::string, int> col, nsid;
    for (int c = 0; c < R; ++c) col[pq.columns[c]] = c;
    for (size_t i = 0; i < pq.namespaces.size(); ++i) nsid[pq.namespaces[i]] = (int)i;
    auto column = [&](const std::string& r) {
        auto it = col.find(r);
        if (it == col.end()) throw Error("resource " + r + " has no column in the packed quotas");
        return it->second;
    };
    auto ns_of = [&](const std::string& n) { auto it = nsid.find(n); return it == nsid.end() ? -1 : it->second; };
    // dense row of a plain resource list, the gpu-memory key skipped; scalar key mask on the side
    auto row = [&](const ResourceList& rl, int64_t* out, bool skipPods, uint32_t* mask) {
        for (auto& [r, v] : rl) {
            if (r == kGpuMemory || (skipPods && r == "pods")) continue;
            const int c = column(r);
            out[c] += v;
            if (mask && c >= 4) *mask |= 1u << c;
        }
    };
    PackedPreemptQuery q;
    q.preemptorNs = ns_of(preemptor.ns); q.preemptorPriority = preemptor.priority;
    q.nominatedInEQ.assign(R, 0); q.nominated.assign(R, 0); q.podReq.assign(R, 0); q.fitRequest.assign(R, 0);
    for (auto& [r, v] : nominatedInEQ) q.nominatedInEQ[column(r)] += v;
    for (auto& [r, v] : nominated) q.nominated[column(r)] += v;
    row(preemptor.request, q.podReq.data(), false, nullptr);           // quota request: gpu-memory filled (pkg/gpu/util/resource.go:60-86)
    int64_t g = 0;
    for (int c = 0; c < R; ++c) g += pq.gpuMemoryCoeff[c] * q.podReq[c];
    q.podReq[pq.gpuMemoryCol] = g;
    row(preemptor.request, q.fitRequest.data(), true, &q.fitMask);     // NodeResourcesFit request
    q.nodePtr.push_back(0); q.podPdbPtr.push_back(0);
    q.nodeAlloc.assign(nodes.size() * R, 0); q.nodeRequested.assign(nodes.size() * R, 0);
    q.pdbAllowed = pdbAllowed;
    for (size_t k = 0; k < nodes.size(); ++k) {
        const PreemptNode& n = nodes[k];
        row(n.allocatable, &q.nodeAlloc[k * R], false, nullptr);
        for (size_t i = 0; i < n.pods.size(); ++i) {
            const QuotaPod& p = n.pods[i];
            q.ns.push_back(ns_of(p.ns)); q.priority.push_back(p.priority); q.overQuota.push_back(p.overQuota ? 1 : 0);
            q.request.resize(q.request.size() + R, 0);
            uint32_t m = 0;
            row(p.request, &q.request[q.request.size() - R], false, &m);
            q.requestMask.push_back(m);
            row(p.request, &q.nodeRequested[k * R], true, nullptr);
            if (i < n.pdbsOfPod.size()) q.podPdb.insert(q.podPdb.end(), n.pdbsOfPod[i].begin(), n.pdbsOfPod[i].end());
            q.podPdbPtr.push_back((int32_t)q.podPdb.size());
        }
        q.nodeRequested[k * R + 3] = (int64_t)n.pods.size();
        q.nodePtr.push_back((int32_t)q.ns.size());
    }
    return q;
}
}  // namespace capacityscheduling

namespace core {

const Node* Snapshot::GetNode(const std::string& name) const {
    auto it = nodes_.find(name);
    return it == nodes_.end() ? nullptr : &it->second;
}

static std::string resource_of_profile(Kind kind, const std::string& p) { return (kind == Kind::MIG ? kMigPrefix : kMpsPrefix) + p; }

state::PartitioningState Snapshot::GetPartitioningState() const {
    // PartitionCalculator.GetPartitioning (internal/partitioning/mig/partitition_calculator.go:30-46, mps twin)
    state::PartitioningState s;
    for (auto& [name, n] : nodes_) {
        state::NodePartitioning np;
        for (auto& g : n.GPUs) {
            state::GPUPartitioning gp; gp.GPUIndex = g.index;
            for (auto& [p, q] : g.GetGeometry()) gp.Resources[resource_of_profile(kind_, p)] = q;
            np.GPUs.push_back(std::move(gp));
        }
        s[name] = std::move(np);
    }
    return s;
}

Snapshot TakeSnapshot(Kind kind, const state::ClusterState& cluster) {
    std::map<std::string, Node> nodes;
    const std::string want = kind == Kind::MIG ? "mig" : "mps";
    for (auto& [k, info] : cluster) {
        if (!info.node) continue;
        auto l = info.node->labels.find(kLabelPartitioning);                // gpu.IsMigPartitioningEnabled / IsMpsPartitioningEnabled
        if (l == info.node->labels.end() || l->second != want) continue;
        nodes.emplace(k, kind == Kind::MIG ? mig::NewNode(info) : slicing::NewNode(info));
    }
    return Snapshot(kind, std::move(nodes));
}

PartitioningPlan NewPartitioningPlan(state::PartitioningState s) {
    auto now = std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count();
    return PartitioningPlan{std::move(s), std::to_string(now)};
}

// ---- packing ---------------------------------------------------------------------------------------------------
Packed Pack(const Snapshot& snapshot, const std::vector<v1::Pod>& pods) {
    const Kind kind = snapshot.kind();
    const auto& nodes = snapshot.GetNodes();                                 // std::map: byte-wise name order (snapshot.go:126-128)
    const auto& known = mig::GetKnownGeometries();
    Packed pk;

    std::vector<ResourceList> reqs; reqs.reserve(pods.size());
    for (auto& p : pods) reqs.push_back(resource::ComputePodRequest(p));

    // interning: a resource name is classified once
    struct Cls { int kind; std::string profile; };                           // 0 cpu 1 mem 2 eph 3 pods 4 slice 5 other
    std::unordered_map<std::string, Cls> cls;
    std::set<std::string> profSet, otherSet;
    auto classify = [&](const std::string& r) -> const Cls& {
        auto it = cls.find(r);
        if (it != cls.end()) return it->second;
        Cls c{5, {}};
        if (r == "cpu") c.kind = 0; else if (r == "memory") c.kind = 1; else if (r == "ephemeral-storage") c.kind = 2;
        else if (r == "pods") c.kind = 3;
        else if (auto p = slice_profile_of(kind, r)) { c.kind = 4; c.profile = *p; profSet.insert(*p); }
        else otherSet.insert(r);
        return cls.emplace(r, std::move(c)).first->second;
    };
    for (auto& [name, n] : nodes) {
        for (auto& g : n.GPUs) {
            for (auto& [p, q] : g.used) profSet.insert(p);
            for (auto& [p, q] : g.free) profSet.insert(p);
            if (std::find(pk.models.begin(), pk.models.end(), g.model) == pk.models.end()) pk.models.push_back(g.model);
        }
        if (n.nodeInfo.node) for (auto& [r, v] : n.nodeInfo.node->allocatable) classify(r);
        for (auto& [r, v] : n.nodeInfo.requested) classify(r);
    }
    for (auto& rq : reqs) for (auto& [r, v] : rq) classify(r);
    if (kind == Kind::MIG) {
        for (auto& m : pk.models) {
            auto it = known.find(m);
            if (it == known.end()) throw Error("model " + m + " is not associated with any known GPU");
            for (auto& geom : it->second) for (auto& [p, q] : geom) profSet.insert(p);
        }
    } else {
        for (auto& p : profSet)
            if (p != std::to_string(mps_size_gb(p)) + "gb")
                throw Error("non-canonical MPS profile name \"" + p + "\" (slicing.NewProfile writes \"%dgb\")");
    }
    pk.profiles.assign(profSet.begin(), profSet.end());                      // MIG: ascending name
    if (kind == Kind::MPS)
        std::sort(pk.profiles.begin(), pk.profiles.end(), [](const std::string& a, const std::string& b) {
            int sa = mps_size_gb(a), sb = mps_size_gb(b);
            return sa != sb ? sa < sb : a < b;
        });
    pk.others.assign(otherSet.begin(), otherSet.end());
    if (pk.profiles.size() > NOSPLAN_MAX_PROFILES) throw Error(std::to_string(pk.profiles.size()) + " distinct profiles > 16");
    if (pk.others.size() > NOSPLAN_MAX_OTHER) throw Error(std::to_string(pk.others.size()) + " non-slice scalar resources > 12");
    const int K = (int)pk.profiles.size(), R = pk.R();
    std::unordered_map<std::string, int> pid, oid;
    for (int i = 0; i < K; i++) pid[pk.profiles[i]] = i;
    for (int i = 0; i < (int)pk.others.size(); i++) oid[pk.others[i]] = 3 + K + i;
    for (auto& p : pk.profiles) {
        pk.profileMemGB.push_back(kind == Kind::MIG ? mig_mem_gb(p) : mps_size_gb(p));
        pk.profileGi.push_back(kind == Kind::MIG ? mig_gi(p) : 0);
    }
    auto col = [&](const std::string& r) -> int {                            // -1 = the pods count, not a column
        const Cls& c = classify(r);
        if (c.kind <= 2) return c.kind;
        if (c.kind == 3) return -1;
        return c.kind == 4 ? 3 + pid.at(c.profile) : oid.at(r);
    };

    // geometry tables
    pk.modelGeomOffset.push_back(0);
    for (auto& m : pk.models) {
        if (kind == Kind::MIG) {
            auto& list = known.at(m);
            if (list.size() > NOSPLAN_MAX_GEOMS) throw Error("model " + m + ": too many geometries");
            for (auto& geom : list) {
                size_t at = pk.geomCounts.size();
                pk.geomCounts.resize(at + 16, 0);
                for (auto& [p, q] : geom) {
                    if (q < 1 || q > 255) throw Error("geometry quantity out of range");
                    pk.geomCounts[at + pid.at(p)] = (uint8_t)q;
                }
            }
        }
        pk.modelGeomOffset.push_back((int32_t)(pk.geomCounts.size() / 16));
    }

    // nodes
    const int N = (int)nodes.size();
    int G = 1;
    for (auto& [name, n] : nodes) G = std::max<int>(G, (int)n.GPUs.size());
    if (G > NOSPLAN_MAX_GPUS) throw Error(std::to_string(G) + " GPUs per node > 16");
    pk.G = G;
    pk.nodeGpuCount.assign(N, 0);
    pk.gpuModel.assign((size_t)N * G, 0); pk.gpuIndex.assign((size_t)N * G, 0); pk.gpuMemoryGB.assign((size_t)N * G, 0);
    pk.gpuUsed.assign((size_t)N * G * 16, 0); pk.gpuFree.assign((size_t)N * G * 16, 0);
    pk.gpuUsedMask.assign((size_t)N * G, 0); pk.gpuFreeMask.assign((size_t)N * G, 0);
    pk.nodeAlloc.assign((size_t)N * R, 0); pk.nodeRequested.assign((size_t)N * R, 0);
    pk.nodeAllowedPods.assign(N, 0); pk.nodeNumPods.assign(N, 0);
    int i = 0;
    for (auto& [name, n] : nodes) {
        pk.nodeNames.push_back(name);
        pk.nodeGpuCount[i] = (int32_t)n.GPUs.size();
        for (size_t j = 0; j < n.GPUs.size(); j++) {
            const GPU& g = n.GPUs[j];
            size_t x = (size_t)i * G + j;
            pk.gpuModel[x] = (int32_t)(std::find(pk.models.begin(), pk.models.end(), g.model) - pk.models.begin());
            pk.gpuIndex[x] = g.index; pk.gpuMemoryGB[x] = g.memoryGB;
            for (auto& [p, q] : g.used) {
                if (q < 0 || q > 255) throw Error("device count out of range");
                pk.gpuUsed[x * 16 + pid.at(p)] = (uint8_t)q; pk.gpuUsedMask[x] |= (uint16_t)(1u << pid.at(p));
            }
            for (auto& [p, q] : g.free) {
                if (q < 0 || q > 255) throw Error("device count out of range");
                pk.gpuFree[x * 16 + pid.at(p)] = (uint8_t)q; pk.gpuFreeMask[x] |= (uint16_t)(1u << pid.at(p));
            }
        }
        if (n.nodeInfo.node)
            for (auto& [r, v] : n.nodeInfo.node->allocatable) {
                int c = col(r);
                if (c < 0) pk.nodeAllowedPods[i] += (int32_t)v; else pk.nodeAlloc[(size_t)i * R + c] += v;
            }
        for (auto& [r, v] : n.nodeInfo.requested) { int c = col(r); if (c >= 0) pk.nodeRequested[(size_t)i * R + c] += v; }
        pk.nodeNumPods[i] = n.nodeInfo.numPods;
        i++;
    }

    // pods, in the caller's order (the engine applies NewPodSorter)
    const int P = (int)pods.size();
    std::unordered_map<std::string, int> keyId;
    pk.podBase.assign((size_t)P * 3, 0);
    pk.scalarPtr.assign(P + 1, 0);
    for (int k = 0; k < P; k++) {
        pk.podPriority.push_back(pods[k].priority.value_or(0));              // corev1.PodPriority
        std::string key = pods[k].ns + "/" + pods[k].name;                   // core/tracker.go:38
        auto it = keyId.find(key);
        if (it == keyId.end()) { it = keyId.emplace(key, (int)keyId.size()).first; pk.podKeys.push_back(key); }
        pk.podTrackerKey.push_back(it->second);
        std::vector<std::pair<int, int64_t>> ent;
        for (auto& [r, v] : reqs[k]) {
            int c = col(r);
            if (c < 0) continue;                                             // "pods" lands in AllowedPodNumber: never read on this path
            if (c < 3) pk.podBase[(size_t)k * 3 + c] += v; else ent.emplace_back(c, v);
        }
        std::sort(ent.begin(), ent.end());
        for (auto& [c, v] : ent) { pk.scalarCol.push_back(c); pk.scalarQty.push_back(v); }
        pk.scalarPtr[k + 1] = (int32_t)pk.scalarCol.size();
    }
    return pk;
}

nosplan_snapshot Packed::snapshot(Kind kind) const {
    nosplan_snapshot s{};
    s.kind = (int32_t)kind; s.n_profiles = (int32_t)profiles.size();
    s.profile_mem_gb = profileMemGB.data(); s.profile_gi = profileGi.data();
    s.n_models = (int32_t)models.size(); s.model_geom_offset = modelGeomOffset.data(); s.geom_counts = geomCounts.data();
    s.n_nodes = (int32_t)nodeGpuCount.size(); s.gpus_per_node = G; s.node_gpu_count = nodeGpuCount.data();
    s.gpu_model = gpuModel.data(); s.gpu_index = gpuIndex.data(); s.gpu_memory_gb = gpuMemoryGB.data();
    s.gpu_used = gpuUsed.data(); s.gpu_free = gpuFree.data(); s.gpu_used_mask = gpuUsedMask.data(); s.gpu_free_mask = gpuFreeMask.data();
    s.n_other = (int32_t)others.size(); s.node_alloc = nodeAlloc.data(); s.node_requested = nodeRequested.data();
    s.node_allowed_pods = nodeAllowedPods.data(); s.node_num_pods = nodeNumPods.data();
    return s;
}
nosplan_pods Packed::pods() const {
    nosplan_pods p{};
    p.n_pods = (int32_t)podPriority.size(); p.priority = podPriority.data(); p.tracker_key = podTrackerKey.data();
    p.base = podBase.data(); p.scalar_ptr = scalarPtr.data(); p.scalar_col = scalarCol.data(); p.scalar_qty = scalarQty.data();
    return p;
}

// ---- planner ---------------------------------------------------------------------------------------------------
Planner::Planner(Kind kind, SchedulerFramework scheduler, int cudaDevice) : kind_(kind), scheduler_(std::move(scheduler)) {
    int rc = nosplan_create(cudaDevice, &h_);
    if (rc != NOSPLAN_OK) {
        std::string msg = std::string("nosplan_create failed (") + std::to_string(rc) + "): " + nosplan_last_error(h_);
        if (h_) nosplan_destroy(h_);
        h_ = nullptr;
        throw Error(msg);
    }
}
Planner::~Planner() { if (h_) nosplan_destroy(h_); }

PartitioningPlan Planner::Plan(const Snapshot& snapshot, const std::vector<v1::Pod>& pendingPods) {
    if (snapshot.kind() != kind_) throw Error("snapshot kind does not match the planner");
    Packed pk = Pack(snapshot, pendingPods);
    std::vector<uint8_t> pre;
    if (scheduler_.preFilter) { pre.reserve(pendingPods.size()); for (auto& p : pendingPods) pre.push_back(scheduler_.preFilter(p) ? 1 : 0); }
    const size_t NG = pk.gpuModel.size();
    std::vector<uint8_t> cnt(NG * 16, 0);
    std::vector<uint16_t> msk(NG, 0);
    nosplan_snapshot s = pk.snapshot(kind_);
    nosplan_pods p = pk.pods();
    nosplan_partitioning out{cnt.data(), msk.data()};
    int rc = nosplan_plan(h_, &s, &p, pre.empty() ? nullptr : pre.data(), scheduler_.filterMode, &out, &stats_);
    if (rc != NOSPLAN_OK) throw Error(std::string("nosplan_plan failed (") + std::to_string(rc) + "): " + nosplan_last_error(h_));
    state::PartitioningState st;
    for (size_t i = 0; i < pk.nodeNames.size(); i++) {
        state::NodePartitioning np;
        for (int j = 0; j < pk.nodeGpuCount[i]; j++) {
            size_t x = i * pk.G + j;
            state::GPUPartitioning gp; gp.GPUIndex = pk.gpuIndex[x];
            for (size_t k = 0; k < pk.profiles.size(); k++)
                if (msk[x] >> k & 1) gp.Resources[resource_of_profile(kind_, pk.profiles[k])] = cnt[x * 16 + k];
            np.GPUs.push_back(std::move(gp));
        }
        st[pk.nodeNames[i]] = std::move(np);
    }
    return NewPartitioningPlan(std::move(st));
}

std::vector<std::string> NodesToPartition(const Snapshot& snapshot, const PartitioningPlan& plan, bool changedOnly) {
    std::vector<std::string> out;
    state::PartitioningState cur = snapshot.GetPartitioningState();
    if (cur.Equal(plan.DesiredState) || plan.DesiredState.IsEmpty()) return out;
    for (auto& [name, np] : plan.DesiredState) {
        if (changedOnly) { auto it = cur.find(name); if (it != cur.end() && it->second.Equal(np)) continue; }
        out.push_back(name);
    }
    return out;
}

ApplyResult Apply(Kind kind, const state::ClusterState& cluster, const Snapshot& snapshot, const PartitioningPlan& plan) {
    ApplyResult res;
    if (snapshot.GetPartitioningState().Equal(plan.DesiredState) || plan.DesiredState.IsEmpty()) return res;
    for (auto& [name, np] : plan.DesiredState) {
        auto it = cluster.find(name);
        if (it == cluster.end() || !it->second.node) throw Error("failed to get node " + name + ": not found");
        NodePatch patch;
        patch.node = name;
        try {
            if (kind == Kind::MIG) {
                auto list = mig::GetGPUSpecAnnotationList(np);
                for (auto& [k, v] : it->second.node->annotations)
                    if (!starts_with(k, kSpecPrefix)) patch.annotations[k] = v;
                for (auto& a : list) patch.annotations[mig::SpecAnnotationKey(a)] = std::to_string(a.Quantity);
                patch.annotations["nos.nebuly.com/spec-partitioning-plan"] = plan.GetId();
            } else {
                patch.config = mps::ToPluginConfig(np);
                patch.configKey = name + "-" + plan.GetId();
            }
        } catch (const Error& e) {
            throw Error("error partitioning node " + name + ": " + e.what());
        }
        res.patches.push_back(std::move(patch));
    }
    res.applied = true;
    return res;
}

}  // namespace core
}  // namespace nos
