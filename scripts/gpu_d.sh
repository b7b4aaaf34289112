#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -x -q -k "c3_full or c4_shape or query_batches or tensor_kernel_variants or hard_distributions or chunked or large_k" > gpurun_out/d_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/d_pytest.log; tail -5 gpurun_out/d_pytest.log
for args in "10000000 1024 100 384" "1250000 1024 100 384" "10000000 4096 100 384" "4000000 1024 100 512" "10000000 256 10 128"; do
  timeout 300 python scripts/qs_ab.py $args >> gpurun_out/d_qs_ab.txt 2>&1
done
cat gpurun_out/d_qs_ab.txt | grep -E "N=|Error|error" 
