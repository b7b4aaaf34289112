#!/bin/bash
mkdir -p gpurun_out
for n in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/i8_bench$n.json 2> gpurun_out/i8_bench$n.err
echo "bench$n exit $?"
done
python - <<'PY'
import json
for n in (4,8):
    for line in open('gpurun_out/i8_bench%d.json'%n):
        if line.startswith('{'):
            j=json.loads(line)
            print("N=%d:"%n, round(j["value"]), "ms", round(j["ms_per_step"],3), "e2e", round(j["e2e"]["value"]), "frac", round(j["roofline"]["frac"],3), "launch_ms", round(j["roofline"]["launch_ms"],3), j["parity_check"]["ok"], j["config"].get("exchange","-")[:20], "clk", j["clocks"]["sm_mhz"], j["clocks"]["reasons"], "p50", round(j["p50_nq1_ms"],3), j["p50_nq1_host_ms"], "3x", round(j["roofline_3xtf32"]["frac"],3))
            c=j.get("c4")
            if c: print("   c4:", round(c["qps"]), round(c["ms_per_step"],3), "e2e", round(c["e2e"]["value"]), "frac", round(c["roofline"]["frac"],3), round(c["roofline"]["launch_ms"],3), c["parity_check"]["ok"], "p50", round(c["p50_nq1_ms"],3), round(c["shard_main_pass_qps"]))
PY
B200KNN_EXCHANGE=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29530 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline --rows 10000000 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('N=8 nccl exchange:', round(j['value']), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']), j['config'].get('exchange'))"
