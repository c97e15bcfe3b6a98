// nos_host_cli -- line-oriented driver of the C++ host mirror, used by the tests to run the reference's own table
// tests (tests/golden/*.json) through nos::core::Planner without a Go toolchain.
//
// stdin (tokens separated by blanks; names carry no blanks):
//   kind mig|mps            filter <NOSPLAN_FILTER_*>        prefilter <0|1>   (constant PreFilter verdict)
//   geometry <model> <profile>=<n>,...   (replaces the default table; one line per allowed geometry)
//   node <name> | nilnode <key>          label <k> <v>   annotation <k> <v>   allocatable <res> <n>   requested <res> <n>
//   numpods <n>
//   pod <namespace> <name> <priority|->  container <res>=<n>,...|-   init <res>=<n>,...|-
//   sa|sb <node> [<gpu index> <res>=<n>,...]   (two partitioning states for `run equal`)
//   gpumem <GB>             quota <name> <maxEnforced 0|1>   qns <namespace>   qmin|qmax|qused <res>=<n>,...|nil
//   pre <ns> <name> <priority> <res>=..   nomeq|nom <res>=..   pdb <allowed>   pnode <allocatable>   ppod <ns> <name> <prio> <over> <res>=.. <pdb ids|->
//                           (`run preemptpack`: the preemption query packer)
//   qset                    qpod <namespace> <name> <priority> <creationTs> <res>=<n>,...|-     (`run quotapack`: the quota packer)
//   planid <id>             (plan id for `run apply`; default "plan")
//   clientdrop <node>       (`run apply`: the node is in the snapshot but the API client does not find it)
//   run plan | run pack | run state | run diff | run equal | run apply | run pluginconfig | run parsespec | run quotapack | run quota | run preemptpack | run preempt
// stdout: `run plan`  -> "node <name>" / "gpu <index> <res>=<n>,..." lines, then "stats ..." ;
//         `run pack`  -> one "<array> v v v" line per packed buffer;  `run state` -> the snapshot's current state;
//         `run diff`  -> plan, then "apply <node>" per node Actuator.Apply would partition and "changed <node>".
//         `run apply` -> Actuator.Apply of the desired state `sa` on the cluster: "applied true|false", then per patched node
//                        "node <name>" and "annotation <k> <v>" lines (MIG) or "configkey <k>" and
//                        "resource <name> <rename> <memoryGB> <device> <replicas>" lines (MPS);
//         `run pluginconfig` -> mps.ToPluginConfig of node "n" of `sa`: "resources <count>" and the resource lines.
// Any nos::Error ends the run with "error <message>" and exit status 1.
#include <cstdio>
#include <iostream>
#include <sstream>

#include "nos/host.hpp"

using namespace nos;

static ResourceList parse_list(const std::string& s) {
    ResourceList rl;
    if (s == "-") return rl;
    std::stringstream ss(s);
    std::string item;
    while (std::getline(ss, item, ',')) {
        size_t eq = item.rfind('=');
        if (eq == std::string::npos) throw Error("bad resource item " + item);
        rl[item.substr(0, eq)] = std::stoll(item.substr(eq + 1));
    }
    return rl;
}

template <class T> static void dump(const char* name, const std::vector<T>& v) {
    std::cout << name;
    for (auto& x : v) std::cout << ' ' << (long long)x;
    std::cout << '\n';
}
static void dump(const char* name, const std::vector<std::string>& v) {
    std::cout << name;
    for (auto& x : v) std::cout << ' ' << x;
    std::cout << '\n';
}
static void print_state(const state::PartitioningState& st) {
    for (auto& [name, np] : st) {
        std::cout << "node " << name << '\n';
        for (auto& g : np.GPUs) {
            std::cout << "gpu " << g.GPUIndex << ' ';
            bool first = true;
            for (auto& [r, q] : g.Resources) { std::cout << (first ? "" : ",") << r << '=' << q; first = false; }
            if (first) std::cout << '-';
            std::cout << '\n';
        }
    }
}

int main() {
    core::Kind kind = core::Kind::MIG;
    core::SchedulerFramework fw;
    int constPre = 1;
    state::ClusterState cluster;
    framework::NodeInfo* cur = nullptr;
    std::vector<v1::Pod> pods;
    mig::KnownGeometries custom;
    state::PartitioningState stA, stB;
    std::string line, mode, planId = "plan";
    std::vector<std::string> clientDrop;
    std::vector<capacityscheduling::ElasticQuotaInfo> quotas;
    std::vector<std::vector<capacityscheduling::QuotaPod>> podSets;
    int gpuMem = 0;
    capacityscheduling::QuotaPod preemptor;
    ResourceList nomInEQ, nomAll;
    std::vector<capacityscheduling::PreemptNode> pnodes;
    std::vector<int32_t> pdbAllowed;
    try {
        while (std::getline(std::cin, line)) {
            std::stringstream ss(line);
            std::string cmd, a, b, c;
            if (!(ss >> cmd) || cmd[0] == '#') continue;
            ss >> a >> b >> c;
            if (cmd == "kind") kind = a == "mps" ? core::Kind::MPS : core::Kind::MIG;
            else if (cmd == "filter") fw.filterMode = std::stoi(a);
            else if (cmd == "prefilter") constPre = std::stoi(a);
            else if (cmd == "geometry") { gpu::Geometry g; for (auto& [p, q] : parse_list(b)) g[p] = (int)q; custom[a].push_back(g); }
            else if (cmd == "node") { cur = &cluster[a]; cur->node = v1::Node{}; cur->node->name = a; }
            else if (cmd == "nilnode") { cur = &cluster[a]; cur->node.reset(); }
            else if (cmd == "label") cur->node->labels[a] = b;
            else if (cmd == "annotation") cur->node->annotations[a] = b;
            else if (cmd == "allocatable") cur->node->allocatable[a] = std::stoll(b);
            else if (cmd == "requested") cur->requested[a] = std::stoll(b);
            else if (cmd == "numpods") cur->numPods = std::stoi(a);
            else if (cmd == "pod") { v1::Pod p; p.ns = a; p.name = b; if (c != "-" && !c.empty()) p.priority = std::stoi(c); pods.push_back(p); }
            else if (cmd == "container") pods.back().containers.push_back({parse_list(a)});
            else if (cmd == "init") pods.back().initContainers.push_back({parse_list(a)});
            else if (cmd == "sa" || cmd == "sb") {          // sa|sb <node> [<gpu index> <res>=<n>,...|-]: one GPU of one of two states
                state::PartitioningState& st = cmd == "sa" ? stA : stB;
                state::NodePartitioning& np = st[a];
                if (!b.empty()) {
                    state::GPUPartitioning g; g.GPUIndex = std::stoi(b);
                    for (auto& [r, q] : parse_list(c.empty() ? "-" : c)) g.Resources[r] = (int)q;
                    np.GPUs.push_back(g);
                }
            }
            else if (cmd == "gpumem") gpuMem = std::stoi(a);
            else if (cmd == "quota") { quotas.emplace_back(); quotas.back().ResourceName = a; quotas.back().MaxEnforced = b == "1"; }
            else if (cmd == "qns") quotas.back().Namespaces.push_back(a);
            else if (cmd == "qmin") { if (a != "nil") quotas.back().Min = parse_list(a); }
            else if (cmd == "qmax") { if (a != "nil") quotas.back().Max = parse_list(a); }
            else if (cmd == "qused") { if (a != "nil") quotas.back().Used = parse_list(a); }
            else if (cmd == "qset") podSets.emplace_back();
            else if (cmd == "qpod") {                       // qpod <namespace> <name> <priority> <creation ts> <res>=<n>,...|-
                std::string d, e;
                ss >> d >> e;
                capacityscheduling::QuotaPod qp;
                qp.ns = a; qp.name = b; qp.priority = std::stoi(c); qp.creationTimestamp = std::stoll(d); qp.request = parse_list(e.empty() ? "-" : e);
                podSets.back().push_back(qp);
            }
            else if (cmd == "pre") { std::string d; ss >> d; preemptor.ns = a; preemptor.name = b; preemptor.priority = std::stoi(c); preemptor.request = parse_list(d.empty() ? "-" : d); }
            else if (cmd == "nomeq") nomInEQ = parse_list(a);
            else if (cmd == "nom") nomAll = parse_list(a);
            else if (cmd == "pdb") pdbAllowed.push_back(std::stoi(a));
            else if (cmd == "pnode") { pnodes.emplace_back(); pnodes.back().allocatable = parse_list(a); }
            else if (cmd == "ppod") {                       // ppod <namespace> <name> <priority> <over 0|1> <res>=<n>,...|- <pdb id,...|->
                std::string d, e, f;
                ss >> d >> e >> f;
                capacityscheduling::QuotaPod qp;
                qp.ns = a; qp.name = b; qp.priority = std::stoi(c); qp.overQuota = d == "1"; qp.request = parse_list(e.empty() ? "-" : e);
                std::vector<int32_t> ids;
                if (!f.empty() && f != "-") { std::stringstream fs(f); std::string t; while (std::getline(fs, t, ',')) ids.push_back(std::stoi(t)); }
                pnodes.back().pods.push_back(qp); pnodes.back().pdbsOfPod.push_back(ids);
            }
            else if (cmd == "planid") planId = a;
            else if (cmd == "clientdrop") clientDrop.push_back(a);      // node the API client of `run apply` does not hold
            else if (cmd == "run") { mode = a; break; }
            else throw Error("unknown directive " + cmd);
        }
        if (mode == "equal") { std::cout << "equal " << (stA.Equal(stB) ? "true" : "false") << '\n'; return 0; }
        auto print_resources = [](const mps::PluginConfig& cfg) {
            for (auto& r : cfg.Resources)
                std::cout << "resource " << r.Name << ' ' << r.Rename << ' ' << r.MemoryGB << ' ' << r.Devices[0] << ' ' << r.Replicas << '\n';
        };
        if (mode == "quotapack") {
            capacityscheduling::PackedQuotas pq = capacityscheduling::PackQuotas(quotas, podSets, gpuMem);
            dump("names", pq.names); dump("namespaces", pq.namespaces); dump("columns", pq.columns);
            dump("gpu_memory_col", std::vector<int>{pq.gpuMemoryCol}); dump("ns_quota", pq.nsQuota);
            dump("min", pq.min); dump("max", pq.max); dump("used_init", pq.usedInit);
            dump("min_mask", pq.minMask); dump("max_mask", pq.maxMask); dump("used_mask", pq.usedMask);
            dump("max_enforced", pq.maxEnforced); dump("gpu_memory_coeff", pq.gpuMemoryCoeff);
            for (size_t k = 0; k < pq.podSets.size(); ++k) {
                const auto& pp = pq.podSets[k];
                const std::string t = "set" + std::to_string(k) + "_";
                dump((t + "keys").c_str(), pp.keys); dump((t + "ns").c_str(), pp.ns); dump((t + "request").c_str(), pp.request);
                dump((t + "request_mask").c_str(), pp.requestMask); dump((t + "priority").c_str(), pp.priority);
                dump((t + "creation_ts").c_str(), pp.creationTs); dump((t + "name_rank").c_str(), pp.nameRank);
            }
            return 0;
        }
        if (mode == "quota") {                                  // PackQuotas -> nosplan_quota_recompute (set 0 running, set 1 pending)
            capacityscheduling::PackedQuotas pq = capacityscheduling::PackQuotas(quotas, podSets, gpuMem);
            nosplan_t* h = nullptr;
            if (nosplan_create(0, &h) != NOSPLAN_OK) throw Error("nosplan_create failed (no CUDA device: there is no CPU path)");
            const size_t E = pq.names.size(), R = (size_t)pq.R();
            const size_t Pr = pq.podSets.size() > 0 ? pq.podSets[0].ns.size() : 0, Pp = pq.podSets.size() > 1 ? pq.podSets[1].ns.size() : 0;
            std::vector<int64_t> used(E * R + 1), aggMin(R), aggUsed(R), aggOver(R), guaranteed(E * R + 1), opUsed(E * R + 1);
            std::vector<uint32_t> usedMask(E + 1), aggMasks(3);
            std::vector<uint8_t> prefilter(Pp + 1), overQuota(Pr + 1);
            int32_t ambiguous = 0;
            nosplan_quota_out out{};
            out.used = used.data(); out.used_mask = usedMask.data(); out.agg_min = aggMin.data(); out.agg_used = aggUsed.data();
            out.agg_overquota = aggOver.data(); out.agg_masks = aggMasks.data(); out.guaranteed = guaranteed.data();
            out.prefilter = prefilter.data(); out.over_quota = overQuota.data(); out.operator_used = opUsed.data(); out.ambiguous_order = &ambiguous;
            nosplan_quotas cq = pq.c();
            nosplan_quota_pods cr{}, cp{};
            if (Pr) cr = pq.podSets[0].c();
            if (Pp) cp = pq.podSets[1].c();
            const int rc = nosplan_quota_recompute(h, &cq, Pr ? &cr : nullptr, Pp ? &cp : nullptr, &out, nullptr);
            if (rc != NOSPLAN_OK) { std::string msg = nosplan_last_error(h); nosplan_destroy(h); throw Error("nosplan_quota_recompute: " + msg); }
            nosplan_destroy(h);
            used.resize(E * R); guaranteed.resize(E * R); opUsed.resize(E * R); prefilter.resize(Pp); overQuota.resize(Pr);
            dump("used", used); dump("agg_min", aggMin); dump("agg_used", aggUsed); dump("agg_overquota", aggOver);
            dump("guaranteed", guaranteed); dump("prefilter", prefilter); dump("over_quota", overQuota); dump("operator_used", opUsed);
            return 0;
        }
        if (mode == "parsespec") {                              // gpu.ParseSpecAnnotation on every annotation of every node
            for (auto& [name, ni] : cluster) {
                if (!ni.node) continue;
                for (auto& [k, v] : ni.node->annotations) {
                    auto sa = gpu::ParseSpecAnnotation(k, v);
                    if (sa) std::cout << "spec " << sa->Index << ' ' << sa->ProfileName << ' ' << sa->Quantity << '\n';
                    else std::cout << "notspec " << k << '\n';
                }
            }
            return 0;
        }
        if (mode == "preemptpack" || mode == "preempt") {       // PackQuotas (the preemptor's resources get columns too), then PackPreemptQuery
            auto sets = podSets;
            sets.push_back({preemptor});
            capacityscheduling::PackedQuotas pq = capacityscheduling::PackQuotas(quotas, sets, gpuMem);
            capacityscheduling::PackedPreemptQuery q = capacityscheduling::PackPreemptQuery(pq, preemptor, nomInEQ, nomAll, pnodes, pdbAllowed);
            if (mode == "preempt") {
                // the current Used of every quota comes from the running pods (set 0), then the dry run -- both on the GPU
                nosplan_t* h = nullptr;
                if (nosplan_create(0, &h) != NOSPLAN_OK) throw Error("nosplan_create failed (no CUDA device: there is no CPU path)");
                const size_t E = pq.names.size(), R = (size_t)pq.R();
                std::vector<int64_t> used(E * R + 1);
                std::vector<uint32_t> usedMask(E + 1);
                nosplan_quota_out out{};
                out.used = used.data(); out.used_mask = usedMask.data();
                nosplan_quotas cq = pq.c();
                nosplan_quota_pods cr = pq.podSets[0].c();
                int rc = nosplan_quota_recompute(h, &cq, pq.podSets[0].ns.empty() ? nullptr : &cr, nullptr, &out, nullptr);
                if (rc == NOSPLAN_OK) {
                    cq.used_init = used.data(); cq.used_mask = usedMask.data();
                    nosplan_preempt_query cqy = q.c();
                    std::vector<uint8_t> victim(q.ns.size() + 1);
                    std::vector<int32_t> status(pnodes.size() + 1), viol(pnodes.size() + 1);
                    rc = nosplan_preempt_dryrun(h, &cq, &cqy, victim.data(), status.data(), viol.data(), nullptr);
                    if (rc == NOSPLAN_OK) {
                        victim.resize(q.ns.size()); status.resize(pnodes.size()); viol.resize(pnodes.size());
                        dump("victim", victim); dump("node_status", status); dump("num_violating", viol);
                    }
                }
                std::string msg = rc == NOSPLAN_OK ? "" : nosplan_last_error(h);
                nosplan_destroy(h);
                if (rc != NOSPLAN_OK) throw Error("preemption dry run: " + msg);
                return 0;
            }
            dump("preemptor", std::vector<int>{q.preemptorNs, q.preemptorPriority, (int)q.fitMask});
            dump("nominated_in_eq", q.nominatedInEQ); dump("pod_req", q.podReq); dump("nominated", q.nominated); dump("fit_request", q.fitRequest);
            dump("node_ptr", q.nodePtr); dump("ns", q.ns); dump("priority", q.priority); dump("over_quota", q.overQuota);
            dump("request", q.request); dump("request_mask", q.requestMask); dump("node_alloc", q.nodeAlloc); dump("node_requested", q.nodeRequested);
            dump("pdb_allowed", q.pdbAllowed); dump("pod_pdb_ptr", q.podPdbPtr); dump("pod_pdb", q.podPdb);
            return 0;
        }
        if (mode == "pluginconfig") {
            mps::PluginConfig cfg = mps::ToPluginConfig(stA["n"]);
            std::cout << "resources " << cfg.Resources.size() << " migstrategy " << cfg.MigStrategy << " failgt1 " << cfg.FailRequestsGreaterThanOne << '\n';
            print_resources(cfg);
            return 0;
        }
        if (!custom.empty()) mig::SetKnownGeometries(custom);
        core::Snapshot snap = core::TakeSnapshot(kind, cluster);
        if (mode == "state") { print_state(snap.GetPartitioningState()); return 0; }
        if (mode == "apply") {
            core::PartitioningPlan plan = core::NewPartitioningPlan(stA);
            plan.id = planId;
            state::ClusterState client = cluster;                          // the snapshot was taken; the client may have lost nodes since
            for (auto& n : clientDrop) client.erase(n);
            core::ApplyResult r = core::Apply(kind, client, snap, plan);
            std::cout << "applied " << (r.applied ? "true" : "false") << '\n';
            for (auto& p : r.patches) {
                std::cout << "node " << p.node << '\n';
                for (auto& [k, v] : p.annotations) std::cout << "annotation " << k << ' ' << v << '\n';
                if (kind == core::Kind::MPS) { std::cout << "configkey " << p.configKey << '\n'; print_resources(p.config); }
            }
            return 0;
        }
        if (mode == "pack") {
            core::Packed pk = core::Pack(snap, pods);
            dump("profiles", pk.profiles); dump("models", pk.models); dump("others", pk.others); dump("node_names", pk.nodeNames);
            dump("pod_keys", pk.podKeys);
            dump("profile_mem_gb", pk.profileMemGB); dump("profile_gi", pk.profileGi); dump("model_geom_offset", pk.modelGeomOffset);
            dump("geom_counts", pk.geomCounts); dump("gpus_per_node", std::vector<int>{pk.G});
            dump("node_gpu_count", pk.nodeGpuCount); dump("gpu_model", pk.gpuModel); dump("gpu_index", pk.gpuIndex);
            dump("gpu_memory_gb", pk.gpuMemoryGB); dump("gpu_used", pk.gpuUsed); dump("gpu_free", pk.gpuFree);
            dump("gpu_used_mask", pk.gpuUsedMask); dump("gpu_free_mask", pk.gpuFreeMask);
            dump("node_alloc", pk.nodeAlloc); dump("node_requested", pk.nodeRequested);
            dump("node_allowed_pods", pk.nodeAllowedPods); dump("node_num_pods", pk.nodeNumPods);
            dump("priority", pk.podPriority); dump("tracker_key", pk.podTrackerKey); dump("base", pk.podBase);
            dump("scalar_ptr", pk.scalarPtr); dump("scalar_col", pk.scalarCol); dump("scalar_qty", pk.scalarQty);
            return 0;
        }
        if (mode == "plan" || mode == "diff") {
            if (!constPre) fw.preFilter = [](const v1::Pod&) { return false; };
            core::Planner planner(kind, fw, 0);
            core::PartitioningPlan plan = planner.Plan(snap, pods);
            print_state(plan.DesiredState);
            const nosplan_stats& st = planner.LastStats();
            std::cout << "stats candidates=" << st.candidates << " visited_nodes=" << st.visited_nodes << " placements=" << st.placements
                      << " geometry_evals=" << st.geometry_evals << " candidate_nodes=" << st.candidate_nodes << '\n';
            if (mode == "diff") {
                for (auto& n : core::NodesToPartition(snap, plan, false)) std::cout << "apply " << n << '\n';
                for (auto& n : core::NodesToPartition(snap, plan, true)) std::cout << "changed " << n << '\n';
            }
            return 0;
        }
        throw Error("no run directive");
    } catch (const std::exception& e) {
        std::cout << "error " << e.what() << '\n';
        return 1;
    }
}
