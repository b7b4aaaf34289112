"""Warm-start sample size sweep (keys per query in multiples of k): warm_sweep.py [N] [nq] [k] [d]"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 1_250_000
nq = a[1] if len(a) > 1 else 1024
k = a[2] if len(a) > 2 else 100
d = a[3] if len(a) > 3 else 384
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
ref = None
for wk in (12, 6, 24, 48, 64, 12):
    idx.set_option("tensor_warm_keys", wk)
    step, main = [], []
    for i in range(14):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); D, I = idx.search(q, k); e1.record(); torch.cuda.synchronize()
        if i >= 4:
            step.append(e0.elapsed_time(e1)); main.append(idx.stats()["main_kernel_ms"])
    if ref is None: ref = (D.clone(), I.clone())
    print("N=%d nq=%d k=%d d=%d warm_keys=%d: step %.3f ms (min %.3f), main pass %.3f ms, identical=%s" % (
        N, nq, k, d, wk, statistics.median(step), min(step), statistics.median(main), bool(torch.equal(ref[1], I) and torch.equal(ref[0], D))), flush=True)
