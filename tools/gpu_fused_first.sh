#!/bin/bash
# Run on a GPU box: fused-kernel tests, planner tests, the bench, a launch list and one full ncu capture of k_plan_fused.
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_fused_gpu.py -x -q) > gpurun_out/t_fused.log 2>&1; tail -15 gpurun_out/t_fused.log
(timeout 900 python -m pytest tests/test_plan_gpu.py tests/test_resident_gpu.py tests/test_changed_nodes_gpu.py tests/test_host_mirror.py -x -q) > gpurun_out/t_plan.log 2>&1; tail -6 gpurun_out/t_plan.log
timeout 300 python bench.py --steps 30 > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; tail -c 1500 gpurun_out/bench_fused.json; tail -3 gpurun_out/bench_fused.err
timeout 300 python bench.py --steps 20 --workload mig-100k --no-cpu-baseline --batch 0 > gpurun_out/bench_fused_100k.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/bench_fused_100k.json').read().strip().splitlines()[-1]); print('100k:', d['ms_per_step'], d['kernel_ms'], d['e2e']['ms_per_step'])"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; grep -c k_plan_fused gpurun_out/launches.csv
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_plan_fused -s 3 -c 1 -o gpurun_out/fused_full \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log | cut -c1-200
