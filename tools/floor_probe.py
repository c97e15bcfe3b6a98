"""What a Plan() cannot go below on this box, measured the way bench.py measures `value` (CUDA events on one stream, a 256 MiB
write between iterations to flush L2): (a) one minimal kernel launch, (b) one kernel streaming the 4.08 MB of algorithmic
bytes of the 10k x 8 x 10k workload once (read 2.7 MB of node rows and pods, write 1.4 MB of plan rows)."""
import json
import numpy as np
import torch

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
one = torch.zeros(1, dtype=torch.int32, device=dev)
src = torch.empty(2_640_192 // 4, dtype=torch.int32, device=dev).zero_()
dst = torch.empty(1_440_000 // 4, dtype=torch.int32, device=dev)


def timed(fn, reps=40):
    ms = []
    for k in range(reps + 5):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        if k >= 5:
            ms.append(s.elapsed_time(e))
    return float(np.median(ms)), float(np.percentile(ms, 95))


a = timed(lambda: one.add_(1))
b = timed(lambda: dst.copy_(src[:dst.numel()]))
c = timed(lambda: (dst.copy_(src[:dst.numel()]), one.add_(src.sum(dtype=torch.int32))))
print(json.dumps({"minimal_kernel_us_p50_p95": [1e3 * a[0], 1e3 * a[1]], "copy_1.44MB_us_p50_p95": [1e3 * b[0], 1e3 * b[1]],
                  "read_2.64MB_write_1.44MB_us_p50_p95": [1e3 * c[0], 1e3 * c[1]],
                  "note": "event-timed on one stream behind a 256 MiB L2 flush, like bench.py's `value`; torch elementwise kernels"}))
