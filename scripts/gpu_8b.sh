#!/bin/bash
mkdir -p gpurun_out
for n in 8 4 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/h8_bench$n.json 2> gpurun_out/h8_bench$n.err
echo "bench$n exit $?"
done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/h8_bench1.json 2> gpurun_out/h8_bench1.err; echo "bench1 exit $?"
python - <<'PY'
import json
for n in (1,2,4,8):
    for line in open('gpurun_out/h8_bench%d.json'%n):
        if line.startswith('{'):
            j=json.loads(line)
            print("N=%d:"%n, round(j["value"]), "ms", round(j["ms_per_step"],3), "e2e", round(j["e2e"]["value"]), "frac", round(j["roofline"]["frac"],3), "launch_ms", round(j["roofline"]["launch_ms"],3), j["parity_check"]["ok"], j["config"].get("exchange","-")[:20], "clk", j["clocks"]["sm_mhz"], j["clocks"]["reasons"], j["clocks"].get("power_w"))
            c=j.get("c4")
            if c: print("   c4:", round(c["qps"]), round(c["ms_per_step"],3), "e2e", round(c["e2e"]["value"]), "frac", round(c["roofline"]["frac"],3), round(c["roofline"]["launch_ms"],3), c["parity_check"]["ok"], "p50", round(c["p50_nq1_ms"],3))
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 2>/dev/null | cut -c1-100,1000-1400
timeout 300 python scripts/multi_lat.py 0,1,2,3,4,5,6,7 2>&1 | tail -4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 scripts/shard_lat.py 10000000 1024 100 384 2>&1 | grep "N="
