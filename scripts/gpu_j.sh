#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/j_pytest.log; tail -6 gpurun_out/j_pytest.log
timeout 200 python scripts/seg_sweep.py 1250000 1024 100 384 2>&1 | grep "N=" | head -3
timeout 200 python scripts/seg_sweep.py 10000000 1024 100 384 2>&1 | grep "N=" | head -3
timeout 200 python scripts/seg_sweep.py 10000000 256 10 128 2>&1 | grep "N=" | head -2
timeout 200 python scripts/seg_sweep.py 12500000 256 100 768 2>&1 | grep "N=" | head -2
