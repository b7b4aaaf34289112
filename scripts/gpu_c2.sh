#!/bin/bash
# 2-GPU call: the real two-device tests, the NCCL sharded path, bench --gpus 2 with its parity check
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/c2_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_sharded.py -m gpu -x -q --durations=8 > gpurun_out/c2_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c2_pytest.log
tail -15 gpurun_out/c2_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c2_bench2.json 2> gpurun_out/c2_bench2.err
echo "bench2 exit $?"; tail -3 gpurun_out/c2_bench2.err; cut -c1-1500 gpurun_out/c2_bench2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/shard_lat.py 10000000 1024 100 384 > gpurun_out/c2_shard_lat.txt 2>&1; grep "N=" gpurun_out/c2_shard_lat.txt
timeout 300 python scripts/multi_lat.py 0,1 > gpurun_out/c2_multi_lat.txt 2>&1; tail -8 gpurun_out/c2_multi_lat.txt
