#!/bin/bash
# Run on a multi-GPU box: the multi-GPU tests, then bench.py at every power of two up to the GPU count (weak), and the strong-scaling sweep.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
(timeout 1500 python -m pytest tests/test_multi_gpu.py -x -q) > gpurun_out/multi_tests.log 2>&1; tail -8 gpurun_out/multi_tests.log
for n in 1 2 4 8; do
  [ "$NG" -ge "$n" ] || continue
  if [ $n -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 --batch 0 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
    timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 --batch 0 --no-cpu-baseline --workload mig-100k --scaling strong > gpurun_out/strong_n1.json 2> gpurun_out/strong_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n bench.py --gpus $n --steps 20 --warmup 3 \
        > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 20 --warmup 3 \
        --workload mig-100k --scaling strong > gpurun_out/strong_n$n.json 2> gpurun_out/strong_n$n.err
  fi
  for f in bench_n$n strong_n$n; do
    python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
    print('$f', 'ms', round(d['ms_per_step'],4), 'mean', round(d['timing']['plan_mean_ms'],4), 'max', round(d['timing']['plan_max_ms'],4), 'e2e', round(d['e2e']['ms_per_step'],3), d['parity'], d['kernel_ms']['path'])
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/$f.err').read()[-1500:])
PY
  done
done
