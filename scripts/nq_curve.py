"""Latency / throughput of IndexFlatL2.search against the batch size (default dispatch, device-resident
queries, CUDA events around the public call).  nq_curve.py [N] [d] [k]"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 10_000_000
d = a[1] if len(a) > 1 else 384
k = a[2] if len(a) > 2 else 100
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0)
print("| nq | path | ms per call | queries/s |\n|---|---|---|---|")
for nq in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
    q = b200knn.synth(5678, 0, nq, d)
    ev = []
    for i in range(14):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); idx.search(q, k); e1.record(); torch.cuda.synchronize()
        if i >= 4: ev.append(e0.elapsed_time(e1))
    st = idx.stats()
    ms = statistics.median(ev)
    print("| %d | %s | %.3f | %.0f |" % (nq, st["path"] + (" (" + (st.get("tensor_mode") or st.get("scan_rows")) + ")"), ms, nq / ms * 1e3), flush=True)
