"""Host logic of nosplan_plan, no GPU: the pod structure the single-launch kernel reads (nos_b200/csrc/api.cu, fast_pods_plan /
fast_pods_fill, exposed as nosplan_host_stage_pods).  The engine builds it in two passes over the caller's arrays without ever
sorting; here it is rebuilt independently from the ORACLE's NewPodSorter order (core/util.go:34-73) and compared section by
section: runs, padded positions, (cpu, memory) pairs in common power-of-two units, minima per 8 / 64 / 512 positions, suffix
minima, cumulative sums and eligibility masks, runs per profile."""
import numpy as np
import pytest

from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol

F_NONE = 0x7fffffff


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def rebuild(snap, pods, pre, meta):
    """The blob's pod-side sections from the oracle's order.  -> dict of arrays."""
    order = ol.sort_pods(snap, pods)
    P, cs, ms = pods.n_pods, meta["cpu_shift"], meta["mem_shift"]
    prof = np.array([int(pods.scalar_col[pods.scalar_ptr[i]]) - 3 if pods.scalar_ptr[i + 1] > pods.scalar_ptr[i] else -1 for i in range(P)])
    runs, start = [], 0
    for k in range(1, P + 1):
        if k == P or prof[order[k]] != prof[order[start]] or pods.priority[order[k]] != pods.priority[order[start]]:
            runs.append((start, k)); start = k
    # a run of the structure = maximal stretch of identical requests; equal priority is not required by the format, but the
    # engine's buckets are (priority, profile), so adjacent buckets with the same profile stay separate runs
    pad = sum(-(-(b - a) // 64) * 64 for a, b in runs)
    pods_sec = np.full((pad, 2), F_NONE, np.int64)
    recs, pos = [], 0
    for a, b in runs:
        n = b - a
        for k in range(n):
            i = order[a + k]
            if pre is None or pre[i]:
                pods_sec[pos + k] = (pods.base[i, 0] >> cs, pods.base[i, 1] >> ms)
        recs.append(dict(pos0=pos, nblk=-(-n // 64), n_real=n, p=int(prof[order[a]]), q=1 if prof[order[a]] >= 0 else 0))
        pos += -(-n // 64) * 64
    g8 = pods_sec.reshape(-1, 8, 2)
    l1 = g8.min(axis=1)
    l2 = l1.reshape(-1, 8, 2).min(axis=1)
    n3 = -(-pad // 512)
    l3 = np.full((n3, 2), F_NONE, np.int64)
    for x in range(pad // 64):
        l3[x // 8] = np.minimum(l3[x // 8], l2[x])
    elig = (g8[:, :, 0] != F_NONE)
    sfx = np.zeros_like(l1); cum = np.zeros((pad // 8, 4), np.int64)
    for r in recs:
        x0, x1 = r["pos0"] // 8, r["pos0"] // 8 + r["nblk"] * 8
        sc = sm = cnt = 0
        for x in range(x0, x1):
            sc += int(g8[x, elig[x], 0].sum()); sm += int(g8[x, elig[x], 1].sum()); cnt += int(elig[x].sum())
            cum[x] = (min(sc, F_NONE), min(sm, F_NONE), cnt, int((elig[x] * (1 << np.arange(8))).sum()))
        t = np.array([F_NONE, F_NONE])
        for x in range(x1 - 1, x0 - 1, -1):
            t = np.minimum(t, l1[x]); sfx[x] = t
        r["min_c"], r["min_m"] = (int(v) for v in l1[x0:x1].min(axis=0))
    return dict(pad=pad, pods=pods_sec, l1=l1, l2=l2, l3=l3, sfx=sfx, cum=cum, runs=recs)


def compare(snap, pods, pre, what):
    blob, meta = capi.host_stage_pods(snap, pods, pre)
    assert blob is not None, f"{what}: the two-pass path did not take these pods"
    exp = rebuild(snap, pods, pre, meta)
    assert meta["pad_pods"] == exp["pad"] and meta["n_runs"] == len(exp["runs"]), what
    sec = lambda name, nxt, cols: blob[meta[name]:meta[nxt]].view(np.int32).reshape(-1, cols)
    got_runs = sec("runs", "runs_per_profile", 8)[:meta["n_runs"]]
    for r, e in zip(got_runs, exp["runs"]):
        assert tuple(int(v) for v in r[:7]) == (e["pos0"], e["nblk"], e["n_real"], e["p"], e["q"], e["min_c"], e["min_m"]), (what, r, e)
    pad = exp["pad"]
    assert np.array_equal(sec("pods", "end", 2)[:pad], exp["pods"]), f"{what}: pod pairs"
    assert np.array_equal(sec("l1", "sfx", 2)[:pad // 8], exp["l1"]), f"{what}: minima per 8"
    assert np.array_equal(sec("l2", "l1", 2)[:pad // 64], exp["l2"]), f"{what}: minima per 64"
    assert np.array_equal(sec("l3", "l2", 2)[:len(exp["l3"])], exp["l3"]), f"{what}: minima per 512"
    assert np.array_equal(sec("sfx", "cum", 2)[:pad // 8], exp["sfx"]), f"{what}: suffix minima"
    assert np.array_equal(sec("cum", "pods", 4)[:pad // 8], exp["cum"]), f"{what}: cumulative sums"
    per_prof = blob[meta["runs_per_profile"]:meta["l3"]].view(np.uint64)
    want = np.zeros(17, np.uint64)
    for k, e in enumerate(exp["runs"]):
        want[e["p"] if e["p"] >= 0 else 16] |= np.uint64(1) << np.uint64(k)
    assert np.array_equal(per_prof[:17], want), f"{what}: runs per profile"
    rset = 0
    for e in exp["runs"]:
        if e["p"] >= 0:
            rset |= 1 << e["p"]
    assert meta["rset"] == rset
    return meta


@pytest.mark.parametrize("kind", ["mig", "mps"])
@pytest.mark.parametrize("n_pods", [1, 7, 64, 65, 1000, 10000])
def test_staged_structure_is_the_oracles_order(kind, n_pods):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    snap, pods = gen(50, 8, n_pods, seed=0x51 + n_pods)
    meta = compare(snap, pods, None, f"{kind} {n_pods}")
    assert meta["n_runs"] >= 1


def test_priorities_prefilter_and_units():
    rng = np.random.default_rng(8)
    for seed in range(8):
        n_pods = int(rng.integers(50, 3000))
        snap, pods = synth.mig_cluster(20, 8, n_pods, seed=0x70 + seed)
        pods.priority[:] = rng.choice([-5, 0, 7, 100, 1000, 2_000_000_000, -2_000_000_000], n_pods)
        pre = (rng.random(n_pods) < 0.6).astype(np.uint8)
        if seed % 2:
            pods.base[:, 0] = rng.integers(1, 4001, n_pods)            # no common power of two: unit 1
            pods.base[:, 1] = rng.integers(1, 1 << 27, n_pods)
        meta = compare(snap, pods, pre, f"seed {seed}")
        if seed % 2:
            assert meta["cpu_shift"] <= 3 and meta["mem_shift"] <= 3
        else:
            assert meta["mem_shift"] >= 20                              # the synthetic requests are whole MiB at least


def test_pods_the_two_pass_path_declines():
    snap, pods = synth.mig_cluster(20, 8, 300, seed=3)
    ok = lambda p: capi.host_stage_pods(snap, p)[0] is not None
    assert ok(pods)
    import copy
    p = copy.deepcopy(pods); p.scalar_qty[0] = 2                       # two devices: the GPU choice shows in the plan
    assert not ok(p)
    p = copy.deepcopy(pods); p.tracker_key[5] = p.tracker_key[6]       # two pods share a tracker entry
    assert not ok(p)
    p = copy.deepcopy(pods); p.base[3, 2] = 1 << 20                    # ephemeral storage
    assert not ok(p)
    p = copy.deepcopy(pods); p.base[4, 1] = 1 << 62                    # a cost beyond 2^28 units
    assert not ok(p)
    p = copy.deepcopy(pods); p.priority[:] = np.arange(p.n_pods)       # more priority levels than runs the kernel holds
    assert not ok(p)
