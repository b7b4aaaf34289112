#!/bin/bash
# round-2 GPU call A: tests, bench, per-phase shard timing
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt; free -g >> gpurun_out/a_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/a_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/a_pytest.log
timeout 600 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench exit $?" >> gpurun_out/a_bench.err
for n in 1250000 2500000 5000000; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 scripts/shard_lat.py $n 1024 100 384 >> gpurun_out/a_shard_lat.txt 2>&1
done
tail -5 gpurun_out/a_pytest.log; cat gpurun_out/a_bench.json | cut -c1-3000; tail -3 gpurun_out/a_bench.err; cat gpurun_out/a_shard_lat.txt | grep "N="
