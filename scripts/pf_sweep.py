"""L2 prefetch distance sweep: pf_sweep.py [N] [nq] [k] [d]"""
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
for probe in (0, 3):
    idx.set_option("tensor_probe", probe)
    for pf in (0, 8, 16, 32, 64, 128, 0):
        idx.set_option("tensor_prefetch_tiles", pf)
        main = []
        for i in range(9):
            D, I = idx.search(q, k); torch.cuda.synchronize()
            if i >= 3: main.append(idx.stats()["main_kernel_ms"])
        same = ""
        if probe == 0:
            if ref is None: ref = (D.clone(), I.clone())
            same = " identical=%s" % bool(torch.equal(ref[0], D) and torch.equal(ref[1], I))
        ms = statistics.median(main)
        print("N=%d nq=%d d=%d probe=%d prefetch=%d: main pass %.3f ms = %.0f TFLOP/s%s" % (N, nq, d, probe, pf, ms, 2.0 * nq * N * d / (ms * 1e-3) / 1e12, same), flush=True)
