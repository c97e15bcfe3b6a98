#!/bin/bash
# Run on a GPU box: plan / host-mirror / sharded tests, then the bench once with NOSPLAN_TRACE=1 (host phases, per-copy times on stderr) and once plain.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_plan_gpu.py tests/test_host_mirror.py tests/test_sharded_gpu.py -x -q 2>&1 | tail -8
NOSPLAN_TRACE=1 timeout 100 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_thr.out 2> gpurun_out/bench_thr.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_thr.out").read().strip().splitlines()[-1])
print("traced run:", d["e2e"]["ms_per_step"], "ms e2e;", d["ms_per_step"], "ms resident")
PY
timeout 100 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_nothr.out 2>/dev/null
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_nothr.out").read().strip().splitlines()[-1])
print("plain run:", d["e2e"]["ms_per_step"], "ms e2e")
PY
