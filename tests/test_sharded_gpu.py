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
