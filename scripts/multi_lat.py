"""Single-process multi-GPU handle: search time on C3 (10M x 384, nq=1024, k=100) and nq=1.  multi_lat.py 0,1[,...] [N]"""
import os, sys, statistics, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import numpy as np, torch
import b200knn
devs = [int(v) for v in sys.argv[1].split(",")]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
d, nq, k = 384, 1024, 100
idx = b200knn.IndexFlatL2(d, devices=devs)
idx.reserve(N); t0 = time.perf_counter(); idx.add_synthetic(N, 1234, 0); print("build %.2f s" % (time.perf_counter() - t0))
q_dev = b200knn.synth(5678, 0, nq, d, device=devs[0])
q_pin = torch.empty(q_dev.shape, dtype=torch.float32, pin_memory=True); q_pin.copy_(q_dev); qh = q_pin.numpy()
D_pin = torch.empty((nq, k), dtype=torch.float32, pin_memory=True).numpy(); I_pin = torch.empty((nq, k), dtype=torch.int64, pin_memory=True).numpy()
for name, fn in (("device tensors", lambda: idx.search(q_dev, k)), ("pinned host", lambda: idx.search(qh, k, D=D_pin, I=I_pin)),
                 ("nq=1 host numpy", lambda: idx.search(qh[:1].copy(), k))):
    ts = []
    for i in range(25):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    print("%s devices=%s N=%d: median %.3f ms (min %.3f)  -> %.1f kQPS" % (name, devs, N, statistics.median(ts[5:]), min(ts), (1 if "nq=1" in name else nq) / statistics.median(ts[5:])))
print(idx.stats())
