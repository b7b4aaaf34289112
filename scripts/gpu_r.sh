#!/bin/bash
mkdir -p gpurun_out
for w in c5 c2 c1; do
timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r_bench_$w.json 2> gpurun_out/r_bench_$w.err; echo "bench $w exit $?"
python - <<PY
import json
for line in open('gpurun_out/r_bench_$w.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("$w qps", round(j["value"],1), "ms", round(j["ms_per_step"],4), "e2e", round(j["e2e"]["value"],1), "roof", j["roofline"]["kernel"], round(j["roofline"]["frac"],3), round(j["roofline"]["launch_ms"],4), j["roofline"].get("scan_rows"), j["parity_check"]["ok"], j["parity_check"]["against"][:40], "p50", round(j["p50_nq1_ms"],4), j.get("p50_nq1_host_ms"), (j.get("roofline_nq1_fp32") or {}).get("frac"), (j.get("cpu_baseline") or {}).get("value"))
PY
done
timeout 300 python bench.py --tensor-mode 2 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('c3 tf32 screen', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['achieved'], j['roofline']['launch_ms'])"
timeout 300 python scripts/nq_curve.py 2>&1 | tail -13
