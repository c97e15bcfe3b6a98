#!/bin/bash
# Run on a GPU box: gate-replay tests and the tests whose expectations changed, every random small case, per-kernel times of the
# rows of the results table that ran as a chain.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gate_replay_gpu.py tests/test_plan_gpu.py tests/test_fused_gpu.py tests/test_resident_gpu.py -q -m gpu) > gpurun_out/gate_tests.log 2>&1; tail -25 gpurun_out/gate_tests.log
timeout 300 python tools/dbg_rand.py 0 3200 2>&1 | tail -12
timeout 300 python tools/dbg_rows.py 2>&1 | tail
