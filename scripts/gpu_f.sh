#!/bin/bash
mkdir -p gpurun_out
python rag-llm-k8s_b200/build.py > /dev/null 2>&1
timeout 200 python scripts/seg_sweep.py 1250000 1024 100 384 > gpurun_out/f_seg.txt 2>&1
timeout 200 python scripts/seg_sweep.py 12500000 256 100 768 >> gpurun_out/f_seg.txt 2>&1
grep "N=" gpurun_out/f_seg.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/f_shard_launches.csv python scripts/shard_one.py 1250000 1024 100 384 3 > gpurun_out/f_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/f_shard_launches.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value'); ui=H.index('Metric Unit')
out=[(r[ki][:60], r[vi], r[ui]) for r in rows[hdr+1:]]
# last search = launches after the last k3_init pair; print the tail
for o in out[-40:]: print(o)
PY
