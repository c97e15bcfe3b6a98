"""Node-sharded Plan() across the GPUs of one box: one process per GPU, one Engine per process, the same snapshot
uploaded on every rank, the node-parallel pass split by node range, ONE all-gather of a few KB (counters + bitmap of
placed tracker keys) on the data path, and an all-gather of the plan rows at the end.

The collective is injectable so the protocol can be driven without NCCL: tests emulate several ranks on one GPU by
concatenating the ranks' blocks, and a gloo world_size-2 test covers the torch.distributed plumbing on CPU.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np


def shard_range(n_nodes: int, rank: int, world: int) -> Tuple[int, int]:
    """Slice of the name-ordered node list owned by `rank` (same rule as nosplan_shard_begin): equal chunks of
    ceil(N/world) nodes, the last possibly short or empty."""
    chunk = (n_nodes + world - 1) // world
    lo = min(n_nodes, rank * chunk)
    return lo, min(n_nodes, lo + chunk)


class _DevArray:
    """Zero-copy view of device memory owned by libnosplan, for torch.as_tensor."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2}


def combine_blocks(blocks: np.ndarray, n_keys: int):
    """Host restatement of k_shard_combine, for tests: blocks [world][bytes] uint8 -> (acc[4] uint64, placed bitmap words)."""
    world = blocks.shape[0]
    words = (n_keys + 31) // 32
    acc = np.zeros(4, np.uint64)
    placed = np.zeros(words, np.uint32)
    for r in range(world):
        acc += blocks[r, :32].view(np.uint64)
        placed |= blocks[r, 32:32 + 4 * words].view(np.uint32)
    return acc, placed


class ShardedPlanner:
    """Drives nosplan_shard_begin / all-gather / nosplan_shard_end on one rank.

    all_gather(dst, src): fills dst ([world * len(src)] uint8 device tensor) with every rank's src; default is
    torch.distributed.all_gather_into_tensor on `group`."""

    def __init__(self, engine, rank: int, world: int, device, group=None, all_gather: Optional[Callable] = None):
        import torch
        self.torch = torch
        self.engine, self.rank, self.world, self.device, self.group = engine, rank, world, device, group
        self._all_gather = all_gather or self._dist_all_gather
        self.local = self.gathered = None

    def _dist_all_gather(self, dst, src):
        import torch.distributed as dist
        dist.all_gather_into_tensor(dst, src, group=self.group)

    def prepare(self, n_nodes: int, gpus_per_node: int):
        """Call after engine.upload(): sizes the exchange buffers and maps the engine's device output rows."""
        t = self.torch
        nb = self.engine.shard_exchange_bytes()
        self.nbytes = nb
        self.local = t.zeros(self.nbytes, dtype=t.uint8, device=self.device)
        self.gathered = t.zeros(self.world * self.nbytes, dtype=t.uint8, device=self.device)
        self.N, self.G = n_nodes, gpus_per_node
        self.chunk = (n_nodes + self.world - 1) // self.world
        d_cnt, d_msk = self.engine.device_outputs()
        rows = max(1, n_nodes * gpus_per_node)
        self.out_count = t.as_tensor(_DevArray(d_cnt, (rows, 16), "|u1"), device=self.device)
        self.out_mask = t.as_tensor(_DevArray(d_msk, (rows, 2), "|u1"), device=self.device)    # u16 as byte pairs: NCCL has no int16
        cr = self.chunk * gpus_per_node
        self.even = n_nodes % self.world == 0 and n_nodes > 0        # slices all-gather straight out of the engine's rows
        self.count_all = t.zeros((self.world * cr, 16), dtype=t.uint8, device=self.device)
        self.mask_all = t.zeros((self.world * cr, 2), dtype=t.uint8, device=self.device)
        if not self.even:
            self.count_local = t.zeros((cr, 16), dtype=t.uint8, device=self.device)
            self.mask_local = t.zeros((cr, 2), dtype=t.uint8, device=self.device)

    def run(self, gather_plan: bool = True):
        """One Plan().  Everything is enqueued on the engine's stream; nothing synchronises the host."""
        t = self.torch
        stream = t.cuda.ExternalStream(self.engine.stream(), device=self.device)
        with t.cuda.stream(stream):
            self.engine.shard_begin(self.rank, self.world, self.local.data_ptr(), self.nbytes)
            self._all_gather(self.gathered, self.local)
            self.engine.shard_end(self.gathered.data_ptr(), self.world)
            if gather_plan and self.N > 0:
                lo, hi = shard_range(self.N, self.rank, self.world)
                a, b = lo * self.G, hi * self.G
                if self.even:
                    self._all_gather(self.count_all.view(-1), self.out_count[a:b].reshape(-1))
                    self._all_gather(self.mask_all.view(-1), self.out_mask[a:b].reshape(-1))
                else:
                    self.count_local.zero_(); self.mask_local.zero_()
                    self.count_local[:b - a] = self.out_count[a:b]
                    self.mask_local[:b - a] = self.out_mask[a:b]
                    self._all_gather(self.count_all.view(-1), self.count_local.view(-1))
                    self._all_gather(self.mask_all.view(-1), self.mask_local.view(-1))

    def plan(self):
        """Host copy of the gathered plan -> (gpu_count [N*G,16] uint8, gpu_mask [N*G] uint16).  Valid when the
        node-parallel pass stood (stats path == 1); after a sequential fallback use engine.download() instead."""
        n = self.N * self.G
        return self.count_all[:n].cpu().numpy(), self.mask_all[:n].cpu().numpy().copy().view(np.uint16).reshape(-1)


# ---------------------------------------------------------------------------------------------------------------------
# slice mode: every rank holds ONLY its own slice of the node list
# ---------------------------------------------------------------------------------------------------------------------
_PER_NODE = ("node_gpu_count", "node_alloc", "node_requested", "node_allowed_pods", "node_num_pods")
_PER_GPU = ("gpu_model", "gpu_index", "gpu_memory_gb", "gpu_used", "gpu_free", "gpu_used_mask", "gpu_free_mask")


def slice_snapshot(snap, lo: int, hi: int, pad_to: int = 0):
    """Nodes [lo, hi) of a packed snapshot as a snapshot of its own (same profiles, models and geometry tables), padded
    with GPU-less, resource-less nodes up to `pad_to` nodes so that every rank of a sliced job holds the same count."""
    import copy
    out = copy.copy(snap)
    G = snap.gpus_per_node
    n = max(hi - lo, pad_to)

    def take(a, a_lo, a_hi, rows):
        a = np.asarray(a)
        piece = a[a_lo:a_hi]
        if len(piece) < rows:
            piece = np.concatenate([piece, np.zeros((rows - len(piece),) + a.shape[1:], a.dtype)])
        return np.ascontiguousarray(piece)

    for f in _PER_NODE:
        setattr(out, f, take(getattr(snap, f), lo, hi, n))
    for f in _PER_GPU:
        setattr(out, f, take(getattr(snap, f), lo * G, hi * G, n * G))
    out.node_names = list(snap.node_names[lo:hi]) + [f"~pad-{k}" for k in range(n - (hi - lo))]
    return out


class SlicedPlanner:
    """Drives nosplan_slice_begin / all-gather / nosplan_slice_mid / all-gather / nosplan_shard_end on one rank whose engine
    was given only that rank's slice (slice_snapshot).  all_gather as in ShardedPlanner."""

    def __init__(self, engine, rank: int, world: int, device, group=None, all_gather: Optional[Callable] = None):
        import torch
        self.torch = torch
        self.engine, self.rank, self.world, self.device, self.group = engine, rank, world, device, group
        self._all_gather = all_gather or ShardedPlanner._dist_all_gather.__get__(self)

    def prepare(self):
        """Call after engine.upload()."""
        t = self.torch
        self.nbytes = self.engine.shard_exchange_bytes()
        self.tot_local = t.zeros(256, dtype=t.uint8, device=self.device)
        self.tot_all = t.zeros(self.world * 256, dtype=t.uint8, device=self.device)
        self.local = t.zeros(self.nbytes, dtype=t.uint8, device=self.device)
        self.gathered = t.zeros(self.world * self.nbytes, dtype=t.uint8, device=self.device)

    def run(self):
        t = self.torch
        stream = t.cuda.ExternalStream(self.engine.stream(), device=self.device)
        with t.cuda.stream(stream):
            self.engine.slice_begin(self.tot_local.data_ptr())
            self._all_gather(self.tot_all, self.tot_local)
            self.engine.slice_mid(self.rank, self.world, self.tot_all.data_ptr(), self.local.data_ptr(), self.nbytes)
            self._all_gather(self.gathered, self.local)
            self.engine.shard_end(self.gathered.data_ptr(), self.world)

    def chain(self, send: Optional[Callable] = None, recv: Optional[Callable] = None):
        """For problems the node-parallel pass does not cover (download() raised SlicedNotApplicable on every rank): the
        sequential sweep rank after rank.  Every rank waits for its predecessor's state block, runs the chain over its slice
        and hands its own block on; send(tensor, dst) / recv(tensor, src) default to torch.distributed point-to-point.
        -> this rank's state block (final on the last rank; see chain_counters)."""
        t = self.torch
        import torch.distributed as dist
        n = self.engine.slice_chain_bytes()
        state_in = t.zeros(n, dtype=t.uint8, device=self.device)
        state_out = t.zeros(n, dtype=t.uint8, device=self.device)
        send = send or (lambda x, dst: dist.send(x, dst, group=self.group))
        recv = recv or (lambda x, src: dist.recv(x, src, group=self.group))
        def turn():
            if self.rank > 0:
                recv(state_in, self.rank - 1)
            self.engine.slice_chain(state_in.data_ptr() if self.rank > 0 else 0, state_out.data_ptr())
            if self.rank + 1 < self.world:
                send(state_out, self.rank + 1)

        if str(self.device) == "cpu":                          # host-side plumbing only (tests with a stand-in engine, gloo)
            turn()
            return state_out
        stream = t.cuda.ExternalStream(self.engine.stream(), device=self.device)
        with t.cuda.stream(stream):
            turn()
        stream.synchronize()
        return state_out


def chain_counters(state) -> dict:
    """Counters of a chain state block (ChainHead in nos_b200/csrc/nosplan_dev.cuh) as download()'s stats name them."""
    raw = state.cpu().numpy() if hasattr(state, "cpu") else np.asarray(state)
    lack = raw[:64].view(np.int32)
    c = raw[64 + 128:64 + 128 + 32].view(np.int64)
    return {"visited_nodes": int(c[0]), "geometry_evals": int(c[1]), "placements": int(c[2]), "candidate_nodes": int(c[3]),
            "lacking_at_end": int((lack > 0).sum()), "lacking_at_start": int(raw[64 + 128 + 32 + 4:64 + 128 + 32 + 8].view(np.int32)[0])}


# ------------------------------------------------------------------------------------------------------------
# Quota accounting over several GPUs (collective 4): pods are split between the ranks, every rank
# reserves its own pods into a zero-based Used on its GPU (nosplan_quota_recompute), the partial Used rows are summed and the
# key masks OR-ed across ranks -- int64 sums and bit-ORs, so any reduction order gives the single-GPU result bit for bit --
# and every rank then derives the aggregates, the guaranteed over-quotas and the PreFilter verdicts of ITS pending pods from
# the reduced rows.  The operator's per-quota running sum needs all pods of a quota in one place: `assign_quotas` gives every
# quota one owner rank (balanced by pod count) and `split_pods_by_owner` routes a pod list accordingly; with that split the
# labels and the label changes are rank-local, and the reduction of Used still holds (non-owners contribute zero rows).
# ------------------------------------------------------------------------------------------------------------
def assign_quotas(pq, pod_ns: np.ndarray, world: int) -> np.ndarray:
    """-> owner rank of every quota row [E]: largest quota first onto the least loaded rank (ties: lowest rank, lowest row)."""
    E = len(pq.names)
    e_of = pq.ns_quota[np.asarray(pod_ns)] if len(pod_ns) else np.zeros(0, np.int32)
    counts = np.bincount(e_of[e_of >= 0], minlength=E) if E else np.zeros(0, np.int64)
    owner = np.zeros(E, np.int32)
    load = np.zeros(world, np.int64)
    for e in sorted(range(E), key=lambda e: (-int(counts[e]), e)):
        r = int(np.argmin(load))
        owner[e] = r
        load[r] += int(counts[e])
    return owner


def take_pods(pods, idx: np.ndarray):
    """The sub-list `idx` of a PackedQuotaPods (order kept)."""
    from . import quota_packing as qp
    idx = np.asarray(idx, np.int64)
    return qp.PackedQuotaPods([pods.keys[i] for i in idx] if pods.keys else [], pods.ns[idx], pods.request[idx], pods.request_mask[idx],
                              pods.priority[idx], pods.creation_ts[idx], pods.name_rank[idx])


def split_pods_by_owner(pq, pods, owner: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Indices of the pods rank `rank` holds: those of the quotas it owns; pods of namespaces without a quota go round-robin."""
    e_of = pq.ns_quota[pods.ns] if pods.n_pods else np.zeros(0, np.int32)
    who = np.where(e_of >= 0, owner[np.maximum(e_of, 0)] if len(owner) else 0, np.arange(pods.n_pods) % world)
    return np.flatnonzero(who == rank)


class ShardedQuota:
    """One rank of the sharded quota recomputation.  `all_reduce(array, op)` reduces a host int64 / uint32 array over the ranks
    in place, op in {"sum", "or"}; default: torch.distributed on `device` (NCCL on a GPU, gloo on the CPU)."""

    def __init__(self, engine, rank: int, world: int, device=None, group=None, all_reduce: Optional[Callable] = None):
        self.engine, self.rank, self.world, self.device, self.group = engine, rank, world, device, group
        self.all_reduce = all_reduce or self._dist_all_reduce

    def _dist_all_reduce(self, a: np.ndarray, op: str):
        import torch
        import torch.distributed as dist
        t = torch.from_numpy(a.view(np.int64) if a.dtype == np.int64 else a.astype(np.int64))
        if self.device is not None and str(self.device) != "cpu":
            t = t.to(self.device)
        if op == "sum":
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        else:                                                  # NCCL has no bitwise OR: gather the (tiny) masks and fold them here
            parts = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t, group=self.group)
            for other in parts:
                t = t | other
        a[...] = t.cpu().numpy().astype(a.dtype)

    def recompute(self, pq, running_mine, pending_mine, label: bool = False):
        """running_mine / pending_mine: this rank's share of the pods.  -> dict as Engine.quota_recompute, with `prefilter`
        (and, with label=True and an owner split, `over_quota`) for this rank's pods only; Used, aggregates, guaranteed rows
        and operator_used are the cluster-wide values on every rank."""
        import dataclasses
        E, R = len(pq.names), pq.n_res
        zero = dataclasses.replace(pq, used_init=np.zeros((E, R), np.int64), used_mask=np.zeros(E, np.uint32))
        part, _ = self.engine.quota_recompute(zero, running_mine if (running_mine is not None and running_mine.n_pods) else None, None)
        used = np.ascontiguousarray(part["used"], dtype=np.int64).copy()
        mask = np.ascontiguousarray(part["used_mask"], dtype=np.uint32).copy()
        op_used = np.ascontiguousarray(part["operator_used"], dtype=np.int64).copy() if label else None
        self.all_reduce(used, "sum")
        self.all_reduce(mask, "or")
        full = dataclasses.replace(pq, used_init=pq.used_init + used, used_mask=(pq.used_mask | mask).astype(np.uint32))
        res, st = self.engine.quota_recompute(full, None, pending_mine if (pending_mine is not None and pending_mine.n_pods) else None)
        res = dict(res)
        if label:
            self.all_reduce(op_used, "sum")                    # a quota's row is non-zero on its owner only
            amb = np.array([int(part["ambiguous_order"])], np.int64)
            self.all_reduce(amb, "or")
            res["over_quota"], res["operator_used"], res["ambiguous_order"] = part["over_quota"], op_used, int(amb[0])
        return res, st
