"""Node-sharded Plan(): several ranks emulated on ONE GPU (one Engine per rank, blocks concatenated instead of an NCCL
all-gather -- the ranks run one after the other, nothing waits on a peer) must give the single-rank plan bit for bit."""
import numpy as np
import pytest

from nos_b200 import capi, packing as pk, sharded, synth
from support import oracle_lib as ol

pytestmark = pytest.mark.gpu


def run_emulated(snap, pods, fm, world):
    import torch
    dev = torch.device("cuda", 0)
    engines = [capi.Engine(0) for _ in range(world)]
    planners = []
    for r, e in enumerate(engines):
        e.upload(snap, pods, None, fm)
        planners.append(sharded.ShardedPlanner(e, r, world, dev, all_gather=lambda dst, src: None))
        planners[-1].prepare(snap.n_nodes, snap.gpus_per_node)
    nb = planners[0].nbytes
    for r, (e, p) in enumerate(zip(engines, planners)):                 # phase 1 on every "rank"
        e.shard_begin(r, world, p.local.data_ptr(), nb)
    torch.cuda.synchronize()
    gathered = torch.cat([p.local for p in planners])                   # the all-gather
    stats, rows = [], []
    for r, (e, p) in enumerate(zip(engines, planners)):
        e.shard_end(gathered.data_ptr(), world)
        cnt, msk, st = e.download()
        stats.append(st)
        rows.append((cnt, msk))
    for e in engines:
        e.close()
    return rows, stats, gathered.cpu().numpy().reshape(world, nb)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_equals_single(world):
    for shape, regime in (((203, 8, 1000), "saturating"), ((64, 4, 300), "saturating"), ((150, 8, 700), "satisfiable")):
        snap, pods = synth.mig_cluster(*shape, seed=77 + world, regime=regime)
        fm = pk.FILTER_RESOURCES_FIT
        ocnt, omsk, ost, _ = ol.plan(snap, pods, None, fm, ol.INCREMENTAL)
        rows, stats, blocks = run_emulated(snap, pods, fm, world)
        G = snap.gpus_per_node
        for r, (cnt, msk) in enumerate(rows):
            st = stats[r]
            for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_start", "lacking_at_end"):
                assert st[k] == ost[k], (shape, regime, r, k)
            if st["path"] == 1:                                          # fast path: the rank owns its slice of the plan
                lo, hi = sharded.shard_range(snap.n_nodes, r, world)
                assert np.array_equal(cnt[lo * G:hi * G], ocnt[lo * G:hi * G]) and np.array_equal(msk[lo * G:hi * G], omsk[lo * G:hi * G])
            else:                                                        # sequential fallback: complete on every rank
                assert np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk)
        assert stats[0]["path"] == (1 if regime == "saturating" else 0)
        acc, placed = sharded.combine_blocks(blocks, pods.n_pods)
        if stats[0]["path"] == 1:
            assert int(acc[0]) == ost["placements"] and int(acc[2]) == ost["visited_nodes"]


# ---------------------------------------------------------------------------------------------------------------------
# slice mode: every "rank" uploads only its own slice of the node list
def run_sliced_emulated(snap, pods, fm, world):
    import torch
    dev = torch.device("cuda", 0)
    chunk = (snap.n_nodes + world - 1) // world
    engines, planners, slices = [], [], []
    for r in range(world):
        lo, hi = sharded.shard_range(snap.n_nodes, r, world)
        sl = sharded.slice_snapshot(snap, lo, hi, pad_to=chunk)
        e = capi.Engine(0)
        e.upload(sl, pods, None, fm)
        p = sharded.SlicedPlanner(e, r, world, dev, all_gather=lambda dst, src: None)
        p.prepare()
        engines.append(e); planners.append(p); slices.append((lo, hi))
    for e, p in zip(engines, planners):
        e.slice_begin(p.tot_local.data_ptr())
    torch.cuda.synchronize()
    tot_all = torch.cat([p.tot_local for p in planners])                # first all-gather: totals blocks
    for r, (e, p) in enumerate(zip(engines, planners)):
        e.slice_mid(r, world, tot_all.data_ptr(), p.local.data_ptr(), p.nbytes)
    torch.cuda.synchronize()
    gathered = torch.cat([p.local for p in planners])                   # second all-gather: exchange blocks
    out = []
    for r, e in enumerate(engines):
        e.shard_end(gathered.data_ptr(), world)
        try:
            cnt, msk, st = e.download()
            out.append((cnt, msk, st))
        except capi.SlicedNotApplicable:
            out.append(None)
    chained = None
    if all(o is None for o in out):
        # the sequential sweep rank after rank: each rank starts from the state block its predecessor left
        n = engines[0].slice_chain_bytes()
        states = [torch.zeros(n, dtype=torch.uint8, device=dev) for _ in range(world)]
        chained = []
        for r, e in enumerate(engines):
            e.slice_chain(states[r - 1].data_ptr() if r else 0, states[r].data_ptr())
            cnt, msk, st = e.download()                      # synchronises: the next rank reads a finished block
            chained.append((cnt, msk, st))
        chained = (chained, sharded.chain_counters(states[-1]))
    for e in engines:
        e.close()
    return out, slices, chained


def check_sliced(snap, pods, fm, world):
    """Slice mode against the oracle: same counters and this rank's rows where the node-parallel pass stands (or nothing is
    lacking); where the single-GPU engine needs its sequential sweep, the NOT_APPLICABLE verdict on every rank and then the
    chained sweep (nosplan_slice_chain), rank after rank, with the oracle's rows and counters."""
    ocnt, omsk, ost, _ = ol.plan(snap, pods, None, fm, ol.INCREMENTAL)
    e = capi.Engine(0)
    single_path = e.plan(snap, pods, None, fm)[2]["path"]
    e.close()
    out, slices, chained = run_sliced_emulated(snap, pods, fm, world)
    G = snap.gpus_per_node
    if single_path == 0 and ost["lacking_at_start"] > 0:
        assert all(o is None for o in out) and chained is not None
        rows, final = chained
        for k in ("visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end"):
            assert final[k] == ost[k], (world, k, final[k], ost[k])
        assert rows[-1][2]["candidates"] == ost["candidates"] and rows[-1][2]["path"] == 0
        for (cnt, msk, st), (lo, hi) in zip(rows, slices):
            n = (hi - lo) * G
            assert np.array_equal(cnt[:n], ocnt[lo * G:hi * G]) and np.array_equal(msk[:n], omsk[lo * G:hi * G])
        return "needs-full"
    for (cnt, msk, st), (lo, hi) in zip(out, slices):
        assert st["path"] == single_path
        for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_start", "lacking_at_end"):
            assert st[k] == ost[k], (world, k, st[k], ost[k])
        n = (hi - lo) * G
        assert np.array_equal(cnt[:n], ocnt[lo * G:hi * G]) and np.array_equal(msk[:n], omsk[lo * G:hi * G])
    return "parallel" if single_path == 1 else "nothing-lacking"


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_sliced_equals_single(world, kind):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    seen = set()
    for shape in ((203, 8, 1000), (64, 4, 300), (37, 8, 500)):
        snap, pods = gen(*shape, seed=91 + world, regime="saturating")
        seen.add(check_sliced(snap, pods, pk.FILTER_RESOURCES_FIT, world))
    assert "parallel" in seen


@pytest.mark.parametrize("world", [2, 4])
def test_sliced_chains_the_sequential_sweep_when_it_is_needed(world):
    seen = set()
    for seed in range(5, 11):           # satisfiable demand: unsaturated lacking vector, or nothing lacking at all (planner.go:80-83)
        snap, pods = synth.mig_cluster(150, 8, 700, seed=seed, regime="satisfiable")
        seen.add(check_sliced(snap, pods, pk.FILTER_RESOURCES_FIT, world))
    snap, pods = synth.mig_cluster(60, 8, 2, seed=3, regime="saturating")      # two pods: lacking but far from saturated
    seen.add(check_sliced(snap, pods, pk.FILTER_NONE, world))
    assert "needs-full" in seen


def test_sliced_random_small_cases_through_the_chain():
    """Random small clusters (multi-profile pods, inconsistent node accounts, zero-valued free keys, a PreFilter vector ...):
    whatever route slice mode takes on two or three ranks -- node-parallel pass, nothing lacking, chained sweep -- rows and
    counters are the oracle's."""
    import nos_model as m
    from support import randcases as rc
    from test_plan_gpu import _has_zero_qty_slice
    import torch
    dev = torch.device("cuda", 0)
    routes = {"parallel": 0, "nothing": 0, "chained": 0}
    for seed in range(0, 240):
        kind_name, nodes, pods, pre, fmode = rc.rand_case(seed, max_nodes=7)
        kind = pk.KIND_BY_NAME[kind_name]
        if _has_zero_qty_slice(kind, pods) or len(nodes) < 2:
            continue
        snap, pp = pk.pack(kind, nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
        ocnt, omsk, ost, _ = ol.plan(snap, pp, pre, fmode, ol.INCREMENTAL)
        world = 2 + seed % 2
        G = snap.gpus_per_node
        chunk = (snap.n_nodes + world - 1) // world
        engines, planners, slices = [], [], []
        for r in range(world):
            lo, hi = sharded.shard_range(snap.n_nodes, r, world)
            e = capi.Engine(0)
            e.upload(sharded.slice_snapshot(snap, lo, hi, pad_to=chunk), pp, pre, fmode)
            p = sharded.SlicedPlanner(e, r, world, dev, all_gather=lambda dst, src: None)
            p.prepare()
            engines.append(e); planners.append(p); slices.append((lo, hi))
        for e, p in zip(engines, planners):
            e.slice_begin(p.tot_local.data_ptr())
        torch.cuda.synchronize()
        tot_all = torch.cat([p.tot_local for p in planners])
        for r, (e, p) in enumerate(zip(engines, planners)):
            e.slice_mid(r, world, tot_all.data_ptr(), p.local.data_ptr(), p.nbytes)
        torch.cuda.synchronize()
        gathered = torch.cat([p.local for p in planners])
        rows, need_chain = [], 0
        for e in engines:
            e.shard_end(gathered.data_ptr(), world)
            try:
                rows.append(e.download())
            except capi.SlicedNotApplicable:
                need_chain += 1
        assert need_chain in (0, world), f"seed {seed}: the ranks disagree"
        if need_chain:
            n = engines[0].slice_chain_bytes()
            states = [torch.zeros(n, dtype=torch.uint8, device=dev) for _ in range(world)]
            for r, e in enumerate(engines):
                e.slice_chain(states[r - 1].data_ptr() if r else 0, states[r].data_ptr())
                rows.append(e.download())
            final = sharded.chain_counters(states[-1])
            routes["chained"] += 1
        else:
            final = rows[-1][2]
            routes["parallel" if final["path"] == 1 else "nothing"] += 1
        for e in engines:
            e.close()
        for k in ("visited_nodes", "placements", "geometry_evals", "lacking_at_start", "lacking_at_end"):
            assert final[k] == ost[k], (seed, k, final[k], ost[k])
        for (cnt, msk, st), (lo, hi) in zip(rows, slices):
            nrow = (hi - lo) * G
            assert np.array_equal(cnt[:nrow], ocnt[lo * G:hi * G]) and np.array_equal(msk[:nrow], omsk[lo * G:hi * G]), f"seed {seed} rows of rank slice {lo}:{hi}"
    assert routes["chained"] > 40, routes


# ---- quota accounting split over the ranks: one Engine and one thread per emulated rank, the reductions through a barrier ----
class ThreadReduce:
    def __init__(self, world):
        import threading
        self.world, self.slots, self.bar = world, [None] * world, threading.Barrier(world)

    def for_rank(self, rank):
        def all_reduce(a, op):
            self.slots[rank] = a.copy()
            self.bar.wait()
            tot = self.slots[0].copy()
            for other in self.slots[1:]:
                tot = tot + other if op == "sum" else tot | other
            self.bar.wait()                                   # everyone has read the slots before the next reduction reuses them
            a[...] = tot
        return all_reduce


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_quota_equals_single(world):
    """C4 recipe, pods split by quota owner: Used / aggregates / guaranteed / operator_used identical on every rank to the
    single-GPU engine and the oracle; PreFilter verdicts and labels identical for the rank's own pods."""
    import threading
    pq, run, pend = synth.quota_state(500, 100_000, 10_000, seed=0xC4)
    exp = ol.quota(pq, run, pend)
    owner = sharded.assign_quotas(pq, run.ns, world)
    red = ThreadReduce(world)
    engines = [capi.Engine(0) for _ in range(world)]
    out, errs = [None] * world, []

    def work(r):
        try:
            mine = sharded.split_pods_by_owner(pq, run, owner, r, world)
            pmine = np.arange(r, pend.n_pods, world)
            sq = sharded.ShardedQuota(engines[r], r, world, all_reduce=red.for_rank(r))
            got, _ = sq.recompute(pq, sharded.take_pods(run, mine), sharded.take_pods(pend, pmine), label=True)
            out[r] = (mine, pmine, got)
        except Exception as e:                                # a dead thread would leave the others at the barrier
            errs.append(e)
            red.bar.abort()

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    for e in engines:
        e.close()
    assert not errs, errs
    seen = np.zeros(run.n_pods, int)
    for mine, pmine, got in out:
        for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed", "operator_used"):
            assert np.array_equal(got[k], exp[k]), k
        assert np.array_equal(got["prefilter"], exp["prefilter"][pmine]) and np.array_equal(got["over_quota"], exp["over_quota"][mine])
        seen[mine] += 1
    assert (seen == 1).all()
