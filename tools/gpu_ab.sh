#!/bin/bash
(timeout 600 python -m pytest tests/test_fused_gpu.py tests/test_compact_gpu.py tests/test_resident_gpu.py -x -q) 2>&1 | tail -2
for k in 1 2 3; do
  timeout 200 python bench.py --steps 40 --batch 0 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('10k ms', round(d['ms_per_step']*1e3,2), 'kernel', round(d['kernel_ms']['parallel']*1e3,2), 'frac', round(d['roofline']['frac'],4), d['parity'])"
done
