#!/bin/bash
# Run on a GPU box: the whole GPU suite, the default bench line, the ncu launch list of a short bench and one full capture
# of k_preempt_dryrun.  Outputs under gpurun_out/.
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests -m gpu -x -q) > gpurun_out/gpu_tests.log 2>&1; tail -4 gpurun_out/gpu_tests.log
timeout 300 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 400 gpurun_out/bench_default.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --batch 0 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; tail -2 gpurun_out/ncu_bench.log | cut -c1-300
timeout 60 python tests/tools/preempt_bench.py > gpurun_out/preempt_bench.json && \
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_preempt_dryrun -s 3 -c 1 -o gpurun_out/preempt_full \
    python tests/tools/preempt_bench.py > gpurun_out/preempt_ncu.log 2>&1; tail -2 gpurun_out/preempt_ncu.log
