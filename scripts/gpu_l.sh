#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/l_bench$i.json 2> gpurun_out/l_bench$i.err; echo "bench exit $?"
python - <<PY
import json
for line in open('gpurun_out/l_bench$i.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("qps", j["value"], "ms", j["ms_per_step"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], "launch_ms", j["roofline"]["launch_ms"], "burst frac", j["roofline"]["frac_of_burst_peak"])
        print("clocks", j["clocks"], "p50", j["p50_nq1_ms"], j["p50_nq1_host_ms"], "3x", j["roofline_3xtf32"]["frac"], j["roofline_3xtf32"]["launch_ms"], "launches", j["gpu_launches"])
PY
done
timeout 300 python scripts/nq_curve.py > gpurun_out/l_nq_curve.txt 2>&1; tail -12 gpurun_out/l_nq_curve.txt
