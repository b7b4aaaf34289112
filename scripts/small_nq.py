"""Small query batches: scan (fp32 / bf16 rows) vs the tensor-core path.  small_nq.py N d k nq..."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
N, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
nqs = [int(v) for v in sys.argv[4:]] or [1]
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0)
for nq in nqs:
    q = b200knn.synth(5678, 0, nq, d)
    out = []
    for name, opts in (("scan fp32", {"path": 1, "scan_rows": 1}), ("scan bf16", {"path": 1, "scan_rows": 2}), ("tensor", {"path": 2})):
        for kk, v in opts.items(): idx.set_option(kk, v)
        ev = []
        for i in range(16):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); D, I = idx.search(q, k); e1.record(); torch.cuda.synchronize()
            if i >= 4: ev.append(e0.elapsed_time(e1))
        out.append("%s %.3f" % (name, statistics.median(ev)))
    print("N=%d d=%d k=%d nq=%d: %s ms" % (N, d, k, nq, " | ".join(out)), flush=True)
# host (numpy) inputs through the tensor path, for comparison with the device-tensor calls above
import time
for nq in nqs:
    qh = b200knn.synth(5678, 0, nq, d).cpu().numpy()
    idx.set_option("path", 2)
    ts = []
    for i in range(12):
        t0 = time.perf_counter(); idx.search(qh, k); ts.append((time.perf_counter() - t0) * 1e3)
    idx.set_option("path", 0)
    ta = []
    for i in range(12):
        t0 = time.perf_counter(); idx.search(qh, k); ta.append((time.perf_counter() - t0) * 1e3)
    print("host inputs nq=%d: forced tensor %.3f ms, auto (%s) %.3f ms" % (nq, statistics.median(ts[3:]), idx.stats()["path"], statistics.median(ta[3:])), flush=True)
