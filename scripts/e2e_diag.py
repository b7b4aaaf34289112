import sys, os, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import numpy as np, torch, b200knn
N, d, nq, k = 10_000_000, 384, 1024, 100
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d); qh = q.cpu().numpy()
def med(f, n=8):
    ts = []
    for i in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); f(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    return statistics.median(ts[2:])
print("device tensors :", med(lambda: idx.search(q, k)), "ms ; kernel", idx.stats()["main_kernel_ms"])
print("host buffers   :", med(lambda: idx.search(qh, k)), "ms ; kernel", idx.stats()["main_kernel_ms"])
def manual():
    qd = torch.from_numpy(qh).cuda(); D, I = idx.search(qd, k); return D.cpu().numpy(), I.cpu().numpy()
print("manual copies  :", med(manual), "ms")
