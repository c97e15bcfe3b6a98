#!/bin/bash
for k in 1 2; do
for v in "" "NOSPLAN_FUSED_TILE8=1"; do
  env $v timeout 200 python bench.py --steps 40 --batch 0 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', 'ms', round(d['ms_per_step']*1e3,2), 'kernel', round(d['kernel_ms']['parallel']*1e3,2))"
done; done
