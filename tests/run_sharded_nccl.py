"""Multi-GPU parity check of the node-sharded Plan() over real NCCL (not collected by pytest: launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/run_sharded_nccl.py
Both drivers are exercised -- collectives issued by torch.distributed (nos_b200/sharded.py) and by the library itself
(nosplan_run_sharded) -- and every rank compares the gathered plan with the CPU oracle and prints one line per case."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)

import numpy as np
import torch
import torch.distributed as dist

from nos_b200 import capi, packing as pk, sharded, synth
from support import oracle_lib as ol

KEYS = ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    eng = capi.Engine(local)
    uid = [eng.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    eng.shard_init(rank, world, uid[0])
    ok = True
    fused_seen = 0
    for shape, regime in (((4000, 8, 3000), "saturating"), ((1001, 8, 2000), "saturating"), ((500, 8, 900), "satisfiable"), ((60, 8, 2), "saturating"), ((150, 8, 700), "satisfiable-7")):
        seed = 31
        if "-" in regime:
            regime, seed = regime.split("-")[0], int(regime.split("-")[1])
        snap, pods = synth.mig_cluster(*shape, seed=seed, regime=regime)
        ocnt, omsk, ost, _ = ol.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
        # (a) collectives through torch.distributed
        eng.upload(snap, pods, None, pk.FILTER_RESOURCES_FIT)
        pl = sharded.ShardedPlanner(eng, rank, world, dev)
        pl.prepare(snap.n_nodes, snap.gpus_per_node)
        pl.run(gather_plan=True)
        cnt_l, msk_l, st = eng.download()
        cnt, msk = pl.plan() if st["path"] == 1 else (cnt_l, msk_l)
        same_a = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(st[k] == ost[k] for k in KEYS)
        # (b) collectives inside the library
        eng.upload(snap, pods, None, pk.FILTER_RESOURCES_FIT)
        eng.run_sharded(True)
        cnt, msk, st2 = eng.download_sharded()
        same_b = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(st2[k] == ost[k] for k in KEYS)
        # (c) slice mode: this rank uploads only its own slice; torch collectives, then the library's own
        chunk = (snap.n_nodes + world - 1) // world
        lo, hi = sharded.shard_range(snap.n_nodes, rank, world)
        sl = sharded.slice_snapshot(snap, lo, hi, pad_to=chunk)
        G = snap.gpus_per_node
        same_c, chained = True, False
        for driver in ("torch", "library"):
            eng.upload(sl, pods, None, pk.FILTER_RESOURCES_FIT)
            if driver == "torch":
                sp = sharded.SlicedPlanner(eng, rank, world, dev)
                sp.prepare()
                sp.run()
            else:
                eng.run_sliced()
            try:
                cnt, msk, st3 = eng.download()
                fused_seen += int(driver == "library" and st3["fused"] == 1)
                n = (hi - lo) * G
                same_c &= st3["path"] == st["path"] and all(st3[k] == ost[k] for k in KEYS)
                same_c &= np.array_equal(cnt[:n], ocnt[lo * G:hi * G]) and np.array_equal(msk[:n], omsk[lo * G:hi * G])
                if st3["path"] == 0:                      # nothing lacking: the unchanged state is the plan
                    same_c &= ost["lacking_at_start"] == 0
            except capi.SlicedNotApplicable:
                same_c &= st["path"] == 0 and ost["lacking_at_start"] > 0   # the sequential sweep is needed: same verdict on every rank
                # ... and runs rank after rank, the state block travelling over NCCL point-to-point
                chainer = sharded.SlicedPlanner(eng, rank, world, dev)
                state = chainer.chain()
                cnt, msk, st4 = eng.download()
                n = (hi - lo) * G
                same_c &= np.array_equal(cnt[:n], ocnt[lo * G:hi * G]) and np.array_equal(msk[:n], omsk[lo * G:hi * G])
                dist.broadcast(state, src=world - 1)
                fin = sharded.chain_counters(state)
                same_c &= all(fin[k] == ost[k] for k in ("visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_end"))
                chained = True
        ok &= same_a and same_b and same_c
        print(f"rank {rank}/{world} {shape} {regime}: path={st['path']}/{st2['path']} torch-collectives={'OK' if same_a else 'MISMATCH'} "
              f"library-collectives={'OK' if same_b else 'MISMATCH'} sliced[{eng.shard_transport()}]={'OK' if same_c else 'MISMATCH'}{' (chained sweep)' if chained else ''}", flush=True)
    ok &= fused_seen >= 2            # the saturating cases must have gone through the single-launch sliced plan
    print(f"rank {rank}/{world} single-launch sliced plans: {fused_seen}", flush=True)
    # quota accounting with the pods split between the ranks: Used summed / masks OR-ed over NCCL, everything else rank-local
    pq, run, pend = synth.quota_state(300, 40_000, 4_000, seed=0xC4)
    exp = ol.quota(pq, run, pend)
    owner = sharded.assign_quotas(pq, run.ns, world)
    mine = sharded.split_pods_by_owner(pq, run, owner, rank, world)
    pmine = np.arange(rank, pend.n_pods, world)
    sq = sharded.ShardedQuota(eng, rank, world, dev)
    got, _ = sq.recompute(pq, sharded.take_pods(run, mine), sharded.take_pods(pend, pmine), label=True)
    same_q = all(np.array_equal(got[k], exp[k]) for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed", "operator_used"))
    same_q &= np.array_equal(got["prefilter"], exp["prefilter"][pmine]) and np.array_equal(got["over_quota"], exp["over_quota"][mine])
    ok &= bool(same_q)
    print(f"rank {rank}/{world} quota 300 quotas / 40000 running ({len(mine)} here) / 4000 pending: sharded-quota={'OK' if same_q else 'MISMATCH'}", flush=True)
    # the same reduction inside the library (nosplan_quota_recompute_sharded): pods dealt round-robin, one NCCL all-reduce of Used
    rmine = np.arange(rank, run.n_pods, world)
    got2, st2q = eng.quota_recompute(pq, sharded.take_pods(run, rmine), sharded.take_pods(pend, pmine), sharded=True)
    same_q2 = all(np.array_equal(got2[k], exp[k]) for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed"))
    same_q2 &= np.array_equal(got2["prefilter"], exp["prefilter"][pmine])
    ok &= bool(same_q2)
    print(f"rank {rank}/{world} quota through nosplan_quota_recompute_sharded ({len(rmine)} running pods here): {'OK' if same_q2 else 'MISMATCH'}", flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
