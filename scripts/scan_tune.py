"""Sweep the scan kernel's CTAs/SM on the nq=1 workloads (CUDA-event timing, median of 30)."""
import sys, os, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
for (n, d, k) in [(1_000_000, 384, 10), (10_000_000, 384, 100), (4_000_000, 1024, 5)]:
    idx = b200knn.IndexFlatL2(d); idx.reserve(n); idx.add_synthetic(n, 1234, 0)
    q = b200knn.synth(5678, 0, 1, d)
    for c in (1, 2, 3, 4, 6, 8):
        idx.set_option("scan_ctas_per_sm", c)
        ts = []
        for i in range(40):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); idx.search(q, k); e1.record(); torch.cuda.synchronize()
            if i >= 10: ts.append(e0.elapsed_time(e1))
        t = statistics.median(ts)
        print("n=%d d=%d k=%d ctas/sm=%d  %.4f ms  %.0f GB/s (%.1f%% of 6549)" % (n, d, k, c, t, n * d * 4 / t / 1e6, n * d * 4 / t / 1e6 / 65.494), flush=True)
    del idx
