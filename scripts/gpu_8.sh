#!/bin/bash
# 8-GPU call: the driver's scaling command at N=8 (with the c4 block), the single-process handle on 8 devices, multi tests
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/g8_bench8.json 2> gpurun_out/g8_bench8.err
echo "bench8 exit $?"; tail -3 gpurun_out/g8_bench8.err
python - <<'PY'
import json
for line in open('gpurun_out/g8_bench8.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("c3 x8:", j["value"], j["ms_per_step"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], "launch_ms", j["roofline"]["launch_ms"], j["parity_check"]["ok"], "p50", j["p50_nq1_ms"])
        c=j.get("c4"); 
        if c: print("c4:", c["qps"], c["ms_per_step"], "e2e", c["e2e"]["value"], "frac", c["roofline"]["frac"], c["roofline"]["launch_ms"], c["parity_check"], "p50", c["p50_nq1_ms"], c["shard_main_pass_qps"])
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/g8_bench4.json 2> gpurun_out/g8_bench4.err
python - <<'PY'
import json
for line in open('gpurun_out/g8_bench4.json'):
    if line.startswith('{'):
        j=json.loads(line)
        print("c3 x4:", j["value"], j["ms_per_step"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"], j["parity_check"]["ok"])
PY
timeout 300 python scripts/multi_lat.py 0,1,2,3,4,5,6,7 > gpurun_out/g8_multi_lat.txt 2>&1; tail -5 gpurun_out/g8_multi_lat.txt
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 scripts/shard_lat.py 10000000 1024 100 384 2>&1 | grep "N="
