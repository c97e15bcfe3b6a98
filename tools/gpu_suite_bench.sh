#!/bin/bash
# Run on a GPU box (any number of GPUs): the whole GPU suite, the default bench line, the reference arm.
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/gpu_tests.log 2>&1; tail -6 gpurun_out/gpu_tests.log
timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 600 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
  for n in 2 4 8; do
    [ "$NG" -ge "$n" ] || continue
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n bench.py --gpus $n --steps 20 --warmup 3 \
        > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -c 300 gpurun_out/bench_n$n.json; tail -2 gpurun_out/bench_n$n.err
  done
fi
