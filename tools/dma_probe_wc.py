"""H2D rate of a 3.3 MB buffer the CPU rewrites before every copy: default page-locked memory against write-combined
(cudaHostAllocWriteCombined), and D2H of 1.44 MB into default page-locked memory.  Run on a GPU box."""
import ctypes as C
import numpy as np, torch
rt = C.CDLL("libcudart.so.12")
torch.zeros(1, device="cuda")
st = torch.cuda.Stream()
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]
nbytes = 3326464
d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
src = np.random.default_rng(1).integers(0, 255, nbytes, dtype=np.uint8)
def host(flags):
    p = C.c_void_p()
    assert rt.cudaHostAlloc(C.byref(p), nbytes, flags) == 0
    return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,))
def rate(p, a, rewrite, n=30):
    ts = []
    for _ in range(n):
        if rewrite: np.copyto(a, src)
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record(st); rt.cudaMemcpyAsync(d.data_ptr(), p, nbytes, 1, st.cuda_stream); e.record(st); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    return float(np.median(ts))
import time
for name, flags in (("default", 0), ("portable", 1), ("write-combined", 4)):
    p, a = host(flags)
    t0 = time.perf_counter(); np.copyto(a, src); t_w = (time.perf_counter() - t0) * 1e6
    print(f"{name:15s} cpu write {t_w:7.1f} us   H2D after rewrite {rate(p, a, True):6.1f} us ({nbytes / rate(p, a, True) / 1e3:5.1f} GB/s)   H2D untouched {rate(p, a, False):6.1f} us", flush=True)
