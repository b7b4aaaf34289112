"""p50 latency of small searches (CUDA events and host wall clock)."""
import sys, os, statistics, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import numpy as np, torch, b200knn
for (n, d, nq, k) in [(2048, 384, 1, 5), (64, 1024, 1, 5), (100000, 384, 1, 10), (1000000, 384, 1, 10), (1000000, 384, 8, 10), (1000000, 384, 19, 10)]:
    idx = b200knn.IndexFlatL2(d); idx.add_synthetic(n, 1234, 0)
    q = b200knn.synth(5678, 0, nq, d); qh = q.cpu().numpy()
    ev, wall = [], []
    for i in range(60):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); idx.search(q, k); e1.record(); torch.cuda.synchronize()
        t0 = time.perf_counter(); idx.search(qh, k); t1 = time.perf_counter()
        if i >= 10: ev.append(e0.elapsed_time(e1) * 1e3); wall.append((t1 - t0) * 1e6)
    print("n=%d d=%d nq=%d k=%d  device p50 %.1f us   host-buffer call p50 %.1f us" % (n, d, nq, k, statistics.median(ev), statistics.median(wall)), flush=True)
