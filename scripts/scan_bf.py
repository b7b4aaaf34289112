"""nq < 20 scan: fp32 rows vs bf16 copies + re-rank (kernel and call times).  scan_bf.py N d k [nq...]"""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
N, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
nqs = [int(v) for v in sys.argv[4:]] or [1]
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
for nq in nqs:
    q = b200knn.synth(5678, 0, nq, d)
    res = {}
    for rows in (1, 2):
        idx.set_option("scan_rows", rows)
        ev, kern = [], []
        for i in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); D, I = idx.search(q, k); e1.record(); torch.cuda.synchronize()
            if i >= 5: ev.append(e0.elapsed_time(e1)); kern.append(idx.stats()["main_kernel_ms"])
        res[rows] = (statistics.median(ev), statistics.median(kern), D.cpu(), I.cpu(), idx.stats().get("scan_rows"))
    same = bool((res[1][3] == res[2][3]).all())
    print("N=%d d=%d k=%d nq=%d: fp32 call %.3f ms (scan %.3f) | %s call %.3f ms (scan %.3f) | same ids %s, max |dD| %.3g"
          % (N, d, k, nq, res[1][0], res[1][1], res[2][4], res[2][0], res[2][1], same, float((res[1][2] - res[2][2]).abs().max())), flush=True)
