#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_quota_gpu.py tests/test_label_changes_gpu.py tests/test_quota_state_gpu.py -q -m gpu -x) > gpurun_out/quota_tests.log 2>&1; tail -15 gpurun_out/quota_tests.log
timeout 200 python bench.py --workload quota-c4 --steps 20 > gpurun_out/bench_quota.json 2> gpurun_out/bench_quota.err; tail -3 gpurun_out/bench_quota.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quota.json').read().strip().splitlines()[-1])
print('kernels ms', d['ms_per_step'], 'frac', d['roofline']['frac'], d['parity'], 'e2e', d['e2e']['ms_per_step'], 'launches', d['timing'])
PY
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_quota.csv python bench.py --workload quota-c4 --steps 2 --warmup 1 > /dev/null 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/launches_quota.csv')))
h=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]; H=rows[h]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
seq=[(r[ki][:50], float(r[vi].replace(',',''))/1000) for r in rows[h+2:] if len(r)>vi]
for k,v in seq[-7:]: print(f"{v:8.2f} us  {k}")
PY
