#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multi.py tests/test_gpu_parity.py -m gpu -x -q --durations=5 > gpurun_out/o2_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/o2_pytest.log; tail -12 gpurun_out/o2_pytest.log
for ex in p2p nccl; do
B200KNN_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/shard_lat.py 2500000 1024 100 384 2>&1 | grep -E "N=|rror" | tail -3
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/o2_bench2.json 2> gpurun_out/o2_bench2.err
echo "bench2 exit $?"; tail -2 gpurun_out/o2_bench2.err
python - <<'PY'
import json
for line in open('gpurun_out/o2_bench2.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("c3 x2:", j["value"], j["ms_per_step"], "e2e", j["e2e"]["value"], j["roofline"]["frac"], j["parity_check"]["ok"], j["clocks"])
PY
