#!/bin/bash
# Run on a GPU box: the whole GPU suite, the default bench line and the reference arm, the ncu launch list of a short bench and one
# full capture of k_plan_fused, the results table, the quota bench, the floor probe.  Outputs under gpurun_out/.
mkdir -p gpurun_out
(time timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/gpu_tests.log 2>&1; tail -4 gpurun_out/gpu_tests.log
timeout 300 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 300 gpurun_out/bench_default.json; tail -2 gpurun_out/bench_default.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 200 gpurun_out/bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; grep -c k_plan_fused gpurun_out/launches.csv
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_plan_fused -s 3 -c 1 -o gpurun_out/fused_full \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; tail -1 gpurun_out/ncu_full.log | cut -c1-200
timeout 200 python bench.py --workload quota-c4 --steps 20 > gpurun_out/bench_quota.json 2> gpurun_out/bench_quota.err; tail -c 200 gpurun_out/bench_quota.json
timeout 100 python tools/floor_probe.py > gpurun_out/floor.json 2>/dev/null; cat gpurun_out/floor.json
timeout 1500 python tests/tools/results_table.py > gpurun_out/results.md 2> gpurun_out/results.err; tail -2 gpurun_out/results.err; wc -l gpurun_out/results.md
