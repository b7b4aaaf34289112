#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_screen_bounds.py -m gpu -x -q > gpurun_out/g_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest.log; tail -4 gpurun_out/g_pytest.log
timeout 200 python scripts/seg_sweep.py 1250000 1024 100 384 2>&1 | grep "N=" | head -3
timeout 200 python scripts/seg_sweep.py 10000000 1024 100 384 2>&1 | grep "N=" | head -2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/g_shard_launches.csv python scripts/shard_one.py 1250000 1024 100 384 3 > gpurun_out/g_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/g_shard_launches.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
out=[(r[ki][:50], r[vi]) for r in rows[hdr+1:]]
for o in out[-15:]: print(o)
PY
