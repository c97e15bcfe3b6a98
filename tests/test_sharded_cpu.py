"""Host-side plumbing of the node-sharded Plan() under torch.distributed (gloo, world_size 2, CPU): shard ranges, the
all-gather of the per-rank blocks and their OR / sum combination.  The CUDA engine is replaced by a stand-in that fills
the block the way k_plan_parallel does; the kernels themselves are covered by tests/test_sharded_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nos_b200 import sharded


def test_shard_ranges_cover_once():
    for n in (0, 1, 7, 8, 9, 1000, 10_001):
        for w in (1, 2, 3, 8):
            seen = []
            for r in range(w):
                lo, hi = sharded.shard_range(n, r, w)
                assert 0 <= lo <= hi <= n
                seen += list(range(lo, hi))
            assert seen == list(range(n))


class FakeEngine:
    """Stands in for capi.Engine on CPU: records what the ShardedPlanner passes around."""

    def __init__(self, rank, n_keys, placed_keys, counters):
        self.rank, self.n_keys, self.placed_keys, self.counters = rank, n_keys, placed_keys, counters
        self.ended_with = None

    def shard_exchange_bytes(self):
        return 32 + 8 * (((self.n_keys + 31) // 32 + 1) // 2)


def _worker(rank, world, port, n_keys, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    keys = rng.choice(n_keys, size=n_keys // 3, replace=False)
    eng = FakeEngine(rank, n_keys, keys, [10 + rank, 20 + rank, 30 + rank, 0])
    nb = eng.shard_exchange_bytes()
    block = np.zeros(nb, np.uint8)
    block[:32].view(np.uint64)[:] = eng.counters
    words = block[32:].view(np.uint32)
    for k in keys:
        words[k >> 5] |= np.uint32(1 << (k & 31))
    local = torch.from_numpy(block.copy())
    gathered = torch.zeros(world * nb, dtype=torch.uint8)
    planner = sharded.ShardedPlanner(eng, rank, world, torch.device("cpu"))
    planner._dist_all_gather(gathered, local)                       # the collective the GPU path uses (NCCL there)
    acc, placed = sharded.combine_blocks(gathered.numpy().reshape(world, nb), n_keys)
    q.put((rank, acc.tolist(), placed.tolist(), sorted(int(k) for k in keys)))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_all_gather_and_combine():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    world, n_keys = 2, 1000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_keys, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    union = set(res[0][3]) | set(res[1][3])
    for rank, acc, placed, _ in res:
        assert acc[:3] == [21, 41, 61]                               # sums of the two ranks' counters
        got = {32 * w + b for w, v in enumerate(placed) for b in range(32) if v >> b & 1}
        assert got == union                                          # OR of the bitmaps


# ---- quota accounting split over the ranks (nos_b200/sharded.py:ShardedQuota) ----------------------------------------
class OracleQuotaEngine:
    """Stands in for capi.Engine.quota_recompute on CPU (the CUDA path is covered by tests/test_sharded_gpu.py)."""

    def quota_recompute(self, pq, running=None, pending=None):
        from support import oracle_lib as ol
        return ol.quota(pq, running, pending), {}


def _quota_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nos_b200 import synth
    pq, run, pend = synth.quota_state(60, 4000, 600, seed=17)
    pq.used_init[:] = (np.arange(pq.used_init.size).reshape(pq.used_init.shape) % 5) * 1000
    owner = sharded.assign_quotas(pq, run.ns, world)
    mine = sharded.split_pods_by_owner(pq, run, owner, rank, world)
    pmine = np.arange(rank, pend.n_pods, world)
    sq = sharded.ShardedQuota(OracleQuotaEngine(), rank, world, torch.device("cpu"))
    got, _ = sq.recompute(pq, sharded.take_pods(run, mine), sharded.take_pods(pend, pmine), label=True)
    q.put((rank, mine, pmine, {k: np.asarray(v) for k, v in got.items()}))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_sharded_quota():
    from nos_b200 import synth
    from support import oracle_lib as ol
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_quota_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pq, run, pend = synth.quota_state(60, 4000, 600, seed=17)
    pq.used_init[:] = (np.arange(pq.used_init.size).reshape(pq.used_init.shape) % 5) * 1000
    exp = ol.quota(pq, run, pend)
    seen = np.zeros(run.n_pods, int)
    for rank, mine, pmine, got in res:
        for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed", "operator_used"):
            assert np.array_equal(got[k], exp[k]), (rank, k)
        assert np.array_equal(got["prefilter"], exp["prefilter"][pmine]) and np.array_equal(got["over_quota"], exp["over_quota"][mine])
        seen[mine] += 1
    assert (seen == 1).all()                                         # every running pod on exactly one rank


def test_assign_quotas_balances_and_keeps_a_quota_whole():
    from nos_b200 import synth
    pq, run, _ = synth.quota_state(200, 20_000, 0, seed=3)
    for world in (2, 4, 8):
        owner = sharded.assign_quotas(pq, run.ns, world)
        sizes = []
        for r in range(world):
            idx = sharded.split_pods_by_owner(pq, run, owner, r, world)
            e = pq.ns_quota[run.ns[idx]]
            assert (owner[e[e >= 0]] == r).all()
            sizes.append(len(idx))
        biggest = np.bincount(pq.ns_quota[run.ns][pq.ns_quota[run.ns] >= 0]).max()
        assert sum(sizes) == run.n_pods and max(sizes) <= max(biggest, run.n_pods // world + biggest // 2 + 1)


# ---- the chained sweep of slice mode: the state block travels rank after rank (SlicedPlanner.chain) ---------------------
class FakeChainEngine:
    """Stands in for capi.Engine on CPU: its slice_chain adds this rank's share to the counters of the incoming block the way
    k_plan_sweep does in chain mode (tests/test_sharded_gpu.py covers the kernel itself)."""

    def __init__(self, rank):
        self.rank = rank

    def slice_chain_bytes(self):
        return 224 + 8

    def stream(self):
        return 0

    def slice_chain(self, d_in, d_out):
        import ctypes
        n = self.slice_chain_bytes()
        block = np.zeros(n, np.uint8) if not d_in else np.frombuffer((ctypes.c_uint8 * n).from_address(d_in), np.uint8).copy()
        c = block[64 + 128:64 + 128 + 32].view(np.int64)
        c[0] += 10 + self.rank            # visited nodes
        c[2] += 100 * (self.rank + 1)     # placements
        c[3] += 7                         # candidate nodes
        block[:64].view(np.int32)[0] = 3 - self.rank          # lacking of profile 0 shrinks along the chain
        block[224 + self.rank] = 0xff                         # removed keys of this rank
        ctypes.memmove(d_out, block.ctypes.data, n)


def _chain_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sp = sharded.SlicedPlanner(FakeChainEngine(rank), rank, world, torch.device("cpu"))
    state = sp.chain()
    dist.broadcast(state, src=world - 1)
    q.put((rank, sharded.chain_counters(state), state.numpy()[224:].tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_chained_sweep_hand_over():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_chain_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, fin, removed in res:                                   # the last rank's block, on every rank
        assert fin["visited_nodes"] == 10 + 11 and fin["placements"] == 100 + 200 and fin["candidate_nodes"] == 14
        assert fin["lacking_at_end"] == 1 and removed[:2] == [255, 255]
