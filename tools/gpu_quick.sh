#!/bin/bash
# Run on a GPU box: fused-kernel tests, a short bench, the per-launch durations under ncu.
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_fused_gpu.py tests/test_compact_gpu.py -x -q) 2>&1 | tail -8
timeout 300 python bench.py --steps 30 --batch 0 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -2 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('ms_per_step', d['ms_per_step'], 'kernel', d['kernel_ms']['parallel'], 'frac', d['roofline']['frac'], d['parity'], 'e2e', d['e2e']['ms_per_step'], 'resident', d['e2e_resident_snapshot']['ms_per_step'], 'soa', d.get('e2e_soa',{}).get('ms_per_step'), d['e2e']['h2d_bytes_per_step'])
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_quick.csv \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > /dev/null 2>&1; grep k_plan_fused gpurun_out/launches_quick.csv | cut -d, -f15 | head -12 | tr '\n' ' '
