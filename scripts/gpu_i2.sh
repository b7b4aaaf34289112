#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --c4-rows 24000000 > gpurun_out/i2_bench2.json 2> gpurun_out/i2_bench2.err
echo "bench2 exit $?"; tail -3 gpurun_out/i2_bench2.err
python - <<'PY'
import json
for line in open('gpurun_out/i2_bench2.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("c3 x2:", j["value"], j["ms_per_step"], j["e2e"]["value"], j["roofline"]["frac"], j["parity_check"]["ok"])
        print("c4:", json.dumps(j.get("c4"))[:1800])
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/shard_lat.py 2500000 1024 100 384 > gpurun_out/i2_shard_lat.txt 2>&1; grep "N=" gpurun_out/i2_shard_lat.txt || tail -5 gpurun_out/i2_shard_lat.txt
timeout 300 python scripts/multi_lat.py 0,1 2500000 > gpurun_out/i2_multi_lat.txt 2>&1; tail -5 gpurun_out/i2_multi_lat.txt
