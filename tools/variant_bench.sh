#!/bin/bash
# Rebuild libnosplan.so on the GPU box with different -D settings and print the kernel times of each (scratch tool).
cd "$(dirname "$0")/.."
for v in "$@"; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fno-fast-math --fmad=false -shared -cudart shared $v \
     nos_b200/csrc/{api,prepass,sweep,static_scratch,quota,victims,shard_nccl,annotations}.cu -ldl -o nos_b200/libnosplan.so 2>/dev/null
  cuobjdump --dump-resource-usage nos_b200/libnosplan.so 2>/dev/null | grep -A1 "k_plan_parallel" | grep -o "REG:[0-9]* STACK:[0-9]*"
  for w in mig-10k mps-c3; do
    python bench.py --workload $w --no-cpu-baseline --steps 30 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['config']['workload'], round(d['ms_per_step']*1e3,1), 'us; parallel', round(d['kernel_ms']['parallel']*1e3,1))"
  done
done
