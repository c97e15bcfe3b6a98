"""The reference's TestDryRunPreemption fixtures (tests/golden/dryrun_preemption.json) in the shape the preemption test plumbing takes."""
import json
import os

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cases():
    """-> [(case dict, quotas, running, nodes, preemptor, gb, pdb_allowed, extra)] with the node's pods least important first."""
    out = []
    for c in json.load(open(os.path.join(HERE, "golden", "dryrun_preemption.json")))["cases"]:
        pods = sorted(c["pods"], key=lambda p: p["priority"])              # stable: equals keep the fixture's order
        quotas = {}
        for ns, e in c["quotas"].items():
            acc = {}
            for p in pods:
                if p["namespace"] == ns:
                    for r in ("memory", "cpu"):
                        acc[r] = acc.get(r, 0) + p["request"].get(r, 0)
            used = e["used"] or {}
            init = {r: used.get(r, 0) - acc.get(r, 0) for r in set(used) | set(acc)}
            quotas[ns] = {"namespaces": [ns], "min": e["min"], "max": e["max"], "used": init, "max_enforced": False}
        nodes = [{"allocatable": c["allocatable"], "pods": pods, "pdbs_of_pod": [[] for _ in pods]}]
        out.append((c, quotas, pods, nodes, c["preemptor"], 8, [], {}))
    return out
