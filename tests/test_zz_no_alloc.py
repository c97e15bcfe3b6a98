"""nosplan_lacking_static: the tracker initialisation for callers that cannot allocate device memory (a planner
sharing its GPU with latency-sensitive streams, where cudaMalloc would stall them, or one confined to a fixed
memory reservation).  Collected last on purpose: a failure here never hides the rest of the suite.

It is called through the C-ABI with device buffers the test owns, twice in a row on different problems (the second
smaller, so stale rows of the shared static table would show), and its result is checked against what the existing
entry point nosplan_plan reports for the same snapshot and pods -- the number of lacking profiles after
NewSliceTracker -- and, bit for bit, against the lacking vector recomputed on the host."""
import numpy as np
import pytest

from nos_b200 import capi, packing as pk, synth

pytestmark = pytest.mark.gpu


def host_lacking(snap, pods):
    K = snap.n_profiles
    tot = (snap.node_alloc[:, 3:3 + K] - snap.node_requested[:, 3:3 + K]).sum(0)
    avail = np.maximum(tot, 0)
    lack = np.zeros(16, np.int64)
    entries = {}
    for i in range(pods.n_pods):
        for e in range(int(pods.scalar_ptr[i]), int(pods.scalar_ptr[i + 1])):
            c = int(pods.scalar_col[e]) - 3
            if c >= K:
                continue
            d = int(pods.scalar_qty[e]) - int(avail[c])
            if d > 0:
                lack[c] += d
                entries[int(pods.tracker_key[i])] = True
    return tot, lack, len(entries)


def run_static(lib, torch, snap, pods, n_keys):
    dev = torch.device("cuda", 0)
    tot, lack_ref, entries_ref = host_lacking(snap, pods)
    tot16 = np.zeros(16, np.int64)
    tot16[:len(tot)] = tot
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    d_key, d_ptr = t(pods.tracker_key, np.int32), t(pods.scalar_ptr, np.int32)
    d_col, d_qty = t(pods.scalar_col, np.int32), t(pods.scalar_qty, np.int64)
    d_tot = t(tot16, np.int64)
    d_lack = torch.full((16,), -7, dtype=torch.int32, device=dev)        # garbage: the call must clear its outputs
    d_ent = torch.full((1,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc, msg = capi.lacking_static(lib, 0, stream, pods.n_pods, n_keys, snap.n_profiles, d_key.data_ptr(), d_ptr.data_ptr(),
                                  d_col.data_ptr(), d_qty.data_ptr(), d_tot.data_ptr(), d_lack.data_ptr(), d_ent.data_ptr())
    assert rc == 0, msg
    torch.cuda.synchronize()
    return d_lack.cpu().numpy(), int(d_ent.cpu()[0]), lack_ref, entries_ref


def test_static_scratch_matches_plan_twice_in_a_row():
    import torch
    import __graft_entry__ as g
    g.build()
    lib = capi.load_library()
    eng = capi.Engine(0)
    rng = np.random.default_rng(5)
    problems = [synth.mig_cluster(300, 8, 4000, seed=21), synth.mig_cluster(40, 4, 300, seed=22, regime="satisfiable")]
    problems[0][1].tracker_key[:] = rng.integers(0, 2500, 4000)          # duplicates share one tracker entry
    for snap, pods in problems:                                           # two calls in a row, shared static table
        n_keys = int(pods.tracker_key.max()) + 1
        lack, n_ent, lack_ref, ent_ref = run_static(lib, torch, snap, pods, n_keys)
        _, _, st = eng.plan(snap, pods, None, pk.FILTER_REJECT_ALL)      # existing entry point, same inputs
        assert int((lack > 0).sum()) == st["lacking_at_start"]
        assert np.array_equal(lack.astype(np.int64), lack_ref)
        assert n_ent == ent_ref
    eng.close()


def test_static_scratch_flags_bad_device_data_instead_of_writing_out_of_bounds():
    """A tracker key outside [0, n_keys) sits in device memory where the host cannot see it: the kernel skips the pod and
    returns a negative entries counter; the next (clean) call is unaffected."""
    import torch
    import __graft_entry__ as g
    g.build()
    lib = capi.load_library()
    snap, pods = synth.mig_cluster(20, 4, 200, seed=23)
    n_keys = int(pods.tracker_key.max()) + 1
    good = pods.tracker_key.copy()
    pods.tracker_key[7] = n_keys + 5                                      # beyond the rows this call clears
    pods.tracker_key[11] = 1 << 20                                        # beyond the table
    pods.tracker_key[13] = -3
    _, n_ent, _, _ = run_static(lib, torch, snap, pods, n_keys)
    assert n_ent < 0
    pods.tracker_key[:] = good
    lack, n_ent, lack_ref, ent_ref = run_static(lib, torch, snap, pods, n_keys)
    assert n_ent == ent_ref and np.array_equal(lack.astype(np.int64), lack_ref)


def test_static_scratch_rejects_oversize_without_launching():
    import __graft_entry__ as g
    g.build()
    lib = capi.load_library()
    rc, msg = capi.lacking_static(lib, 0, None, 0, capi.STATIC_MAX_KEYS + 1, 4, None, None, None, None, 1, 1, 1)
    assert rc == -3 and "NOSPLAN_STATIC_MAX_KEYS" in msg
