#!/bin/bash
# Run on a GPU box: ncu launch lists (per-launch durations, cold caches, serialised) of the quota, label-change and preemption benches.
mkdir -p gpurun_out
for t in quota label preempt; do
  timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$t.csv \
      python tests/tools/${t}_bench.py > gpurun_out/ncu_$t.log 2>&1
  grep -c "gpu__time_duration" gpurun_out/launches_$t.csv
done
