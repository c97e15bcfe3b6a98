"""Random small logical snapshots (test plumbing): deliberately includes the awkward regimes --
inconsistent Allocatable vs annotations, zero-valued map keys, multi-profile pods, duplicate pod
names, pods without slice requests, over-committed nodes."""
import random

import nos_model as m
from nos_b200 import packing as pk

MIG_MODELS = list(m.DEFAULT_KNOWN_GEOMETRIES)
MPS_SIZES = [5, 10, 20, 30, 40]


def _split(rng, geom):
    used, free = {}, {}
    for p, q in geom.items():
        u = rng.randint(0, q)
        if u or rng.random() < 0.15:
            used[p] = u
        if q - u or rng.random() < 0.3:
            free[p] = q - u
    return used, free


def rand_mig_gpu(rng, model, index):
    r = rng.random()
    geoms = m.DEFAULT_KNOWN_GEOMETRIES[model]
    if r < 0.15:
        used, free = {}, {}
    elif r < 0.85:
        used, free = _split(rng, rng.choice(geoms))
    else:                                   # not an allowed geometry
        other = rng.choice(MIG_MODELS)
        used, free = _split(rng, rng.choice(m.DEFAULT_KNOWN_GEOMETRIES[other]))
        if rng.random() < 0.5 and free:
            free.pop(next(iter(free)))
    return {"model": model, "index": index, "memory_gb": 0, "used": used, "free": free}


def rand_mps_gpu(rng, index, mem):
    used, free, left = {}, {}, mem
    for _ in range(rng.randint(0, 4)):
        s = rng.choice(MPS_SIZES)
        if s > left:
            continue
        left -= s
        d = used if rng.random() < 0.5 else free
        d[f"{s}gb"] = d.get(f"{s}gb", 0) + 1
    if rng.random() < 0.1:
        free[f"{rng.choice(MPS_SIZES)}gb"] = free.get(f"{rng.choice(MPS_SIZES)}gb", 0)
    return {"model": "NVIDIA-A100-80GB-PCIe", "index": index, "memory_gb": mem, "used": used, "free": free}


def rand_case(seed, kind_name=None, max_nodes=6, max_pods=12):
    rng = random.Random(seed)
    kind_name = kind_name or rng.choice(["mig", "mps"])
    kind = pk.KIND_BY_NAME[kind_name]
    prefix = pk.MIG_PREFIX if kind == pk.KIND_MIG else pk.MPS_PREFIX
    consistent = rng.random() < 0.5
    nodes = []
    for i in range(rng.randint(0, max_nodes)):
        ng = rng.choice([0, 1, 1, 2, 2, 3, 4])
        if kind == pk.KIND_MIG:
            model = rng.choice(MIG_MODELS)
            gpus = [rand_mig_gpu(rng, model if rng.random() < 0.9 else rng.choice(MIG_MODELS), j) for j in range(ng)]
        else:
            mem = rng.choice([16, 24, 40, 48, 80, 82])
            gpus = [rand_mps_gpu(rng, j, mem) for j in range(ng)]
        alloc = {"cpu": rng.choice([2000, 8000, 64000]), "memory": rng.choice([8, 64, 512]) << 30,
                 "pods": rng.choice([0, 1, 3, 110])}
        req = {"cpu": rng.choice([0, 500, 1900, 9000]), "memory": rng.choice([0, 4, 60]) << 30}
        if consistent or rng.random() < 0.5:
            for g in gpus:
                for p, q in g["used"].items():
                    alloc[prefix + p] = alloc.get(prefix + p, 0) + q
                    req[prefix + p] = req.get(prefix + p, 0) + q
                for p, q in g["free"].items():
                    alloc[prefix + p] = alloc.get(prefix + p, 0) + q
        else:
            for g in gpus:
                for p in list(g["used"]) + list(g["free"]):
                    if rng.random() < 0.7:
                        alloc[prefix + p] = rng.randint(0, 9)
                    if rng.random() < 0.3:
                        req[prefix + p] = rng.randint(0, 3)
        if rng.random() < 0.3:
            alloc["nvidia.com/gpu"] = rng.randint(0, 4)
        if rng.random() < 0.2:
            alloc["example.com/foo"] = rng.randint(0, 4)
        nodes.append({"name": f"node-{rng.randint(0, 99):02d}-{i}", "gpus": gpus, "allocatable": alloc,
                      "requested": req, "num_pods": rng.choice([0, 0, 2, 110])})
    if kind == pk.KIND_MIG:
        profs = sorted({p for gs in m.DEFAULT_KNOWN_GEOMETRIES.values() for g in gs for p in g})
    else:
        profs = [f"{s}gb" for s in MPS_SIZES] + ["15gb"]
    focus = rng.sample(profs, k=min(len(profs), rng.randint(1, 4)))
    pods = []
    for i in range(rng.randint(0, max_pods)):
        c = {}
        r = rng.random()
        if r < 0.75:
            c[prefix + rng.choice(focus)] = rng.choice([1, 1, 1, 2, 3])
        elif r < 0.87:
            for p in rng.sample(focus, k=min(len(focus), 2)):
                c[prefix + p] = rng.choice([1, 2])
        if rng.random() < 0.6:
            c["cpu"] = rng.choice([100, 1000, 4000])
        if rng.random() < 0.5:
            c["memory"] = rng.choice([1, 8, 32]) << 30
        if rng.random() < 0.1:
            c["nvidia.com/gpu"] = 1
        if rng.random() < 0.05:
            c["example.com/foo"] = rng.randint(0, 2)
        pod = {"namespace": f"ns-{rng.randint(0, 2)}", "name": f"pd-{i if rng.random() < 0.9 else rng.randint(0, 3)}",
               "containers": [c] if c or rng.random() < 0.5 else []}
        if rng.random() < 0.15 and c:
            pod["containers"].append(dict(c))
        if rng.random() < 0.1:
            pod["init_containers"] = [{"cpu": 6000}]
        if rng.random() < 0.6:
            pod["priority"] = rng.choice([0, 1, 5, 100])
        pods.append(pod)
    pre = [1 if rng.random() < 0.85 else 0 for _ in pods]
    fmode = rng.choice([pk.FILTER_NONE, pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT, pk.FILTER_RESOURCES_FIT, pk.FILTER_REJECT_ALL])
    return kind_name, nodes, pods, pre, fmode


def model_snapshot(kind_name, nodes, known=None):
    """logical nodes -> oracle/nos_model.py object graph"""
    known = known or m.DEFAULT_KNOWN_GEOMETRIES
    out = {}
    for n in nodes:
        ni = m.node_info_from(n)
        if kind_name == "mig":
            gpus = [m.MigGPU(g["model"], g["index"], dict(g["used"]), dict(g["free"]), known[g["model"]]) for g in n["gpus"]]
            out[n["name"]] = m.MigNode(n["name"], gpus, ni)
        else:
            gpus = [m.MpsGPU(g["model"], g["index"], g["memory_gb"], dict(g["used"]), dict(g["free"])) for g in n["gpus"]]
            out[n["name"]] = m.MpsNode(n["name"], gpus, ni)
    return m.ClusterSnapshot(kind_name, out)


def model_plan(kind_name, nodes, pods, pre, fmode):
    snap = model_snapshot(kind_name, nodes)
    mpods = [m.pod_from(p) for p in pods]
    if fmode == pk.FILTER_RESOURCES_FIT:
        return m.plan(snap, mpods, pre, m.FILTER_RESOURCES_FIT)
    return m.plan(snap, mpods, pre, m.FILTER_NONE, fmode == pk.FILTER_NONE)
