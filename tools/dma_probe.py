"""H2D copy rate from different kinds of page-locked host memory (run on a GPU box)."""
import ctypes as C, time, sys, os
import numpy as np, torch
rt = C.CDLL("libcudart.so.12")
dev = torch.device("cuda", 0)
torch.cuda.init(); torch.zeros(1, device=dev)
st = torch.cuda.Stream()
def timed(fn, n=20):
    ts = []
    for _ in range(n):
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record(st); fn(); e.record(st); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    return float(np.median(ts))
for nbytes in (160 << 10, 1280 << 10, 5 << 20):
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    tp = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
    p = C.c_void_p()
    assert rt.cudaMallocHost(C.byref(p), C.c_size_t(nbytes)) == 0
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    a = np.zeros(nbytes, np.uint8)
    treg = torch.from_numpy(a)
    assert rt.cudaHostRegister(C.c_void_p(a.ctypes.data), C.c_size_t(nbytes), 0) == 0
    def c_torch():
        with torch.cuda.stream(st): d.copy_(tp, non_blocking=True)
    def c_raw(src):
        return lambda: rt.cudaMemcpyAsync(C.c_void_p(d.data_ptr()), C.c_void_p(src), nbytes, 1, C.c_void_p(st.cuda_stream))
    print(f"{nbytes>>10:6d} KiB  torch-pinned via torch copy_: {timed(c_torch):7.1f} us   torch-pinned raw memcpyAsync: {timed(c_raw(tp.data_ptr())):7.1f} us"
          f"   cudaMallocHost: {timed(c_raw(p.value)):7.1f} us   cudaHostRegister(numpy): {timed(c_raw(a.ctypes.data)):7.1f} us", flush=True)
# many segments back to back
nb = 1280 << 10
d = torch.empty(8 * nb, dtype=torch.uint8, device=dev)
tps = [torch.zeros(nb, dtype=torch.uint8).pin_memory() for _ in range(8)]
def many():
    for k, t in enumerate(tps):
        rt.cudaMemcpyAsync(C.c_void_p(d.data_ptr() + k * nb), C.c_void_p(t.data_ptr()), nb, 1, C.c_void_p(st.cuda_stream))
print("8 x 1280 KiB torch-pinned back to back:", timed(many), "us")
big = torch.zeros(8 * nb, dtype=torch.uint8).pin_memory()
print("1 x 10 MiB torch-pinned:", timed(lambda: rt.cudaMemcpyAsync(C.c_void_p(d.data_ptr()), C.c_void_p(big.data_ptr()), 8 * nb, 1, C.c_void_p(st.cuda_stream))), "us")
