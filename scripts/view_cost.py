"""What rag.py's per-request read_index costs once the file is resident: a view (b2k_view) + its first search."""
import os, sys, time, statistics, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import numpy as np, torch, b200knn as faiss
d, n = 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
idx = faiss.IndexFlatL2(d); idx.add_synthetic(n, 1234, 0)
p = os.path.join(tempfile.mkdtemp(), "faiss_index")
t0 = time.perf_counter(); faiss.write_index(idx, p); print("write_index %d x %d: %.1f ms" % (n, d, (time.perf_counter() - t0) * 1e3))
faiss.clear_index_cache()
t0 = time.perf_counter(); a = faiss.read_index(p); print("read_index (cold, from the file): %.1f ms" % ((time.perf_counter() - t0) * 1e3))
q = faiss.synth(5678, 0, 1, d).cpu().numpy()
rd, se = [], []
for i in range(30):
    t0 = time.perf_counter(); v = faiss.read_index(p); t1 = time.perf_counter(); D, I = v.search(q, 5); t2 = time.perf_counter()
    rd.append((t1 - t0) * 1e3); se.append((t2 - t1) * 1e3); del v
print("read_index (resident): median %.3f ms; first search on the fresh view (nq=1, k=5): median %.3f ms" % (statistics.median(rd[5:]), statistics.median(se[5:])))
