"""Steady-state (power-capped) A/B of the query-stationary schedule on C3: alternating blocks of 60 searches."""
import os, sys, statistics, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
N, nq, k, d = 10_000_000, 1024, 100, 384
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
for i in range(150): idx.search(q, k)     # heat up
torch.cuda.synchronize()
for rep in range(3):
    for qs in (0, 1):
        idx.set_option("tensor_query_stationary", qs)
        for i in range(5): idx.search(q, k)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(60): idx.search(q, k)
        e1.record(); torch.cuda.synchronize()
        print("rep %d qs=%d: %.3f ms per search (60 back to back), last main pass %.3f ms" % (rep, qs, e0.elapsed_time(e1) / 60, idx.stats()["main_kernel_ms"]), flush=True)
