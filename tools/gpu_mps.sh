#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_plan_gpu.py tests/test_resident_gpu.py tests/test_compact_gpu.py -q -m gpu) > gpurun_out/mps_tests.log 2>&1; tail -30 gpurun_out/mps_tests.log
timeout 300 python tools/dbg_rand.py 0 3200 2>&1 | tail -8
timeout 600 python tests/tools/results_table.py --no-faithful "--only=C3" "--only=MPS half" > gpurun_out/results_mps.md 2> gpurun_out/results_mps.err; tail -3 gpurun_out/results_mps.err; tail -16 gpurun_out/results_mps.md | cut -c1-260
