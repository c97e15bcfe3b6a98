"""Test plumbing for the C++ host mirror (host/): writes the line-oriented problem nos_host_cli reads and parses what it prints."""
import os
import subprocess

from nos_b200 import packing as pk

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CLI = os.path.join(ROOT, "host", "nos_host_cli")


def _rl(d):
    return ",".join(f"{k}={v}" for k, v in d.items()) or "-"


def logical_to_v1(kind_name, nodes, strict=True):
    """Logical node (gpus with used/free maps) -> v1.Node labels + status annotations; None when the node cannot be
    expressed that way (GPUs of different models, or -- strict only -- GPU indices that NewNode would renumber)."""
    out = []
    for n in nodes:
        models = {g["model"] for g in n["gpus"]}
        if len(models) > 1:
            return None
        labels = {pk.LABEL_PARTITIONING: kind_name, pk.LABEL_PRODUCT: models.pop() if models else "NVIDIA-A30",
                  pk.LABEL_COUNT: str(len(n["gpus"]))}
        if kind_name == "mps":
            mem = n["gpus"][0]["memory_gb"] if n["gpus"] else 40
            labels[pk.LABEL_MEMORY] = str(mem * 1000)
        ann = {}
        annotated = [bool(g["used"] or g["free"]) for g in n["gpus"]]
        if strict and any(annotated[i] and not all(annotated[:i]) for i in range(len(annotated))):
            return None                      # un-annotated GPUs are appended after the annotated ones
        for g in n["gpus"]:
            for st in ("used", "free"):
                for p, q in g[st].items():
                    ann[f"{pk.STATUS_PREFIX}-{g['index']}-{p}-{st}"] = str(q)
        out.append({"name": n["name"], "labels": labels, "annotations": ann, "allocatable": n.get("allocatable", {}),
                    "requested": n.get("requested", {}), "num_pods": n.get("num_pods", 0)})
    return out


def to_text(kind_name, v1_nodes, pods, mode, filter_mode=0, prefilter=1, geometries=None):
    L = [f"kind {kind_name}", f"filter {filter_mode}", f"prefilter {int(prefilter)}"]
    for model, geoms in (geometries or {}).items():
        L += [f"geometry {model} {_rl(g)}" for g in geoms]
    for n in v1_nodes:
        L.append(f"node {n['name']}")
        L += [f"label {k} {v}" for k, v in n.get("labels", {}).items()]
        L += [f"annotation {k} {v}" for k, v in n.get("annotations", {}).items()]
        L += [f"allocatable {k} {v}" for k, v in n.get("allocatable", {}).items()]
        L += [f"requested {k} {v}" for k, v in n.get("requested", {}).items()]
        L.append(f"numpods {n.get('num_pods', 0)}")
    for p in pods:
        prio = p.get("priority")
        L.append(f"pod {p['namespace']} {p['name']} {'-' if prio is None else prio}")
        if "request" in p:
            L.append(f"container {_rl(p['request'])}")
        L += [f"container {_rl(c)}" for c in p.get("containers", [])]
        L += [f"init {_rl(c)}" for c in p.get("init_containers", [])]
    L.append(f"run {mode}")
    return "\n".join(L) + "\n"


def run(text):
    r = subprocess.run([CLI], input=text, capture_output=True, text=True, timeout=120)
    return r.returncode, r.stdout


def parse_state(out):
    """-> ({node: [(gpu_index, {res: n})]}, {stat: n}, extra lines)"""
    st, stats, extra, cur = {}, {}, [], None
    for line in out.splitlines():
        t = line.split()
        if t[0] == "node":
            cur = st.setdefault(t[1], [])
        elif t[0] == "gpu":
            res = {} if t[2] == "-" else {kv.rsplit("=", 1)[0]: int(kv.rsplit("=", 1)[1]) for kv in t[2].split(",")}
            cur.append((int(t[1]), res))
        elif t[0] == "stats":
            stats = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in t[1:]}
        else:
            extra.append(t)
    return st, stats, extra


def parse_pack(out):
    d = {}
    for line in out.splitlines():
        t = line.split()
        d[t[0]] = t[1:]
    return d
