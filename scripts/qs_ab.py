"""A/B of the query-stationary schedule: qs_ab.py [N] [nq] [k] [d]  -> main-pass and step time with the option off / on."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 10_000_000
nq = a[1] if len(a) > 1 else 1024
k = a[2] if len(a) > 2 else 100
d = a[3] if len(a) > 3 else 384
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
ref = None
for rep in range(2):
    for qs in (0, 1):
        idx.set_option("tensor_query_stationary", qs)
        step, main = [], []
        for i in range(12):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(); D, I = idx.search(q, k); e1.record(); torch.cuda.synchronize()
            if i >= 4:
                step.append(e0.elapsed_time(e1)); main.append(idx.stats()["main_kernel_ms"])
        if ref is None:
            ref = (D.clone(), I.clone())
        same = bool(torch.equal(ref[1], I) and torch.equal(ref[0], D))
        ms = statistics.median(main)
        print("N=%d nq=%d k=%d d=%d qs=%d: step %.3f ms, main pass %.3f ms = %.0f TFLOP/s, %s, identical=%s" % (
            N, nq, k, d, qs, statistics.median(step), ms, 2.0 * nq * N * d / (ms * 1e-3) / 1e12, idx.stats().get("tensor_mode"), same), flush=True)
