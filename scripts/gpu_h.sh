#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/h_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/h_pytest.log; tail -4 gpurun_out/h_pytest.log
timeout 600 python bench.py > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; echo "bench exit $?"; tail -2 gpurun_out/h_bench.err
python - <<'PY'
import json
for line in open('gpurun_out/h_bench.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print(j["value"], j["ms_per_step"], j["e2e"]["value"], j["roofline"]["frac"], j["roofline"]["launch_ms"])
        print(j["roofline_3xtf32"]); print(j["tf32_peak"]); print(j["parity_check"])
PY
