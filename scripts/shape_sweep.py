"""Batched search over a grid of shapes: call time, main-pass time and bf16-equivalent TFLOP/s, to spot
shapes where the heuristics (segments, CTA pairs, warm start) misbehave.  shape_sweep.py"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
print("| N | d | nq | k | call ms | main pass ms | TFLOP/s (2*nq*N*d / call) | mode | fallbacks |\n|---|---|---|---|---|---|---|---|---|")
for N, d in ((100_000, 384), (1_000_000, 384), (4_000_000, 384), (10_000_000, 384), (1_000_000, 64), (3_000_000, 1024), (2_000_000, 768)):
    idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
    for nq in (32, 200, 1024, 4096):
        q = b200knn.synth(5678, 0, nq, d)
        for k in (1, 10, 100, 1000):
            if N * nq * k > 4e12 and k == 1000 and nq == 4096:
                continue
            ev = []
            for i in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); idx.search(q, k); e1.record(); torch.cuda.synchronize()
                if i >= 2: ev.append(e0.elapsed_time(e1))
            st = idx.stats(); ms = statistics.median(ev)
            print("| %d | %d | %d | %d | %.3f | %.3f | %.0f | %s | %s |" % (N, d, nq, k, ms, st.get("main_kernel_ms") or 0, 2.0 * nq * N * d / (ms * 1e-3) / 1e12,
                                                                  st.get("tensor_mode") or st.get("scan_rows"), st.get("screen_fallbacks")), flush=True)
    del idx
