"""Multi-GPU parity check of the node-sharded Plan() over real NCCL (not collected by pytest: launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/run_sharded_nccl.py
Every rank compares the gathered plan with the CPU oracle and prints one line."""
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


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ok = True
    for shape, regime in (((4000, 8, 3000), "saturating"), ((1001, 8, 2000), "saturating"), ((500, 8, 900), "satisfiable")):
        snap, pods = synth.mig_cluster(*shape, seed=31, regime=regime)
        eng = capi.Engine(local)
        eng.upload(snap, pods, None, pk.FILTER_RESOURCES_FIT)
        pl = sharded.ShardedPlanner(eng, rank, world, dev)
        pl.prepare(snap.n_nodes, snap.gpus_per_node)
        pl.run(gather_plan=True)
        cnt_l, msk_l, st = eng.download()
        ocnt, omsk, ost, _ = ol.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
        if st["path"] == 1:
            cnt, msk = pl.plan()
        else:
            cnt, msk = cnt_l, msk_l
        same = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(
            st[k] == ost[k] for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end"))
        ok &= same
        print(f"rank {rank}/{world} {shape} {regime}: path={st['path']} parity={'OK' if same else 'MISMATCH'}", flush=True)
        eng.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
