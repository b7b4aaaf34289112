#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/t_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/t_pytest.log; tail -3 gpurun_out/t_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err; echo "bench exit $?"
python - <<'PY'
import json
for line in open('gpurun_out/t_bench.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("qps", j["value"], "ms", j["ms_per_step"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], j["clocks"])
        print(sorted(j.keys()))
PY
