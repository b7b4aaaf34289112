#!/bin/bash
mkdir -p gpurun_out
python scripts/shard_one.py 1250000 1024 100 384 3 > /dev/null 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/k_shard_launches.csv python scripts/shard_one.py 1250000 1024 100 384 3 > gpurun_out/k_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/k_shard_launches.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value')
out=[(r[ki][:50], r[vi]) for r in rows[hdr+1:]]
for o in out[-16:]: print(o)
PY
# full captures of the refresh (K4) and re-rank (K6) kernels of the third search
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k4_merge|k6_rerank" -s 10 -c 5 -o gpurun_out/k_k4k6 python scripts/shard_one.py 1250000 1024 100 384 3 > gpurun_out/k_ncu2.log 2>&1
ls -la gpurun_out/k_k4k6.ncu-rep
