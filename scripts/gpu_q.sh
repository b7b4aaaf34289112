#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/q_pytest.log; tail -4 gpurun_out/q_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench exit $?"
python - <<'PY'
import json
for line in open('gpurun_out/q_bench.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("qps", j["value"], "ms", j["ms_per_step"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], "launch_ms", j["roofline"]["launch_ms"], j["clocks"])
        print("p50", j["p50_nq1_ms"], j["p50_nq1_host_ms"], j["roofline_nq1"]["frac"], j["roofline_nq1_fp32"]["frac"], "3x", j["roofline_3xtf32"]["frac"], j["roofline_3xtf32"]["launch_ms"], "launches", j["gpu_launches"], j["cpu_baseline"]["value"])
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 | cut -c1-400
