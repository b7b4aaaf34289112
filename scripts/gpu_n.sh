#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/n_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/n_pytest.log; tail -4 gpurun_out/n_pytest.log
timeout 200 python scripts/seg_sweep.py 1250000 1024 100 384 2>&1 | grep "N=" | head -1
timeout 200 python scripts/seg_sweep.py 10000000 4096 1024 1024 2>&1 | grep "N=" | head -1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/n_shard_launches.csv python scripts/shard_one.py 1250000 1024 100 384 3 > gpurun_out/n_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/n_shard_launches.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
out=[(r[ki][:50], r[vi]) for r in rows[hdr+1:]]
for o in out[-3:]: print(o)
PY
python bench.py --workload c3q1 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('c3q1', j['value'], j['ms_per_step'], j['roofline']['frac'], j['p50_nq1_ms'], j.get('p50_nq1_host_ms'))"
