"""Quick GPU check of the tensor-core path against the oracle (run under `timeout`)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import b200knn
from oracle import oracle as O

cases = [(1000, 64, 128, 10), (5000, 384, 40, 10), (100000, 384, 300, 100), (70000, 100, 1024, 5), (3001, 1024, 129, 37)]
if len(sys.argv) > 1:
    cases = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for (n, d, nq, k) in cases:
    x = O.synth(1234, 0, n, d); q = O.synth(5678, 0, nq, d)
    idx = b200knn.IndexFlatL2(d); idx.add(x); idx.set_option("path", 2); idx.set_option("timing", 1)
    t0 = time.time(); D, I = idx.search(q, k); t1 = time.time()
    Dr, Ir = O.search(x, q, k, O.MODE_EXPAND)
    Dt, It = O.truth_f64(x, q, k)
    same = (I == Ir).mean(); same_t = (I == It).mean()
    rel = np.abs(D - Dt).max() / Dt.max()
    print("n=%d d=%d nq=%d k=%d  ids==oracle %.5f ids==f64 %.5f  max|D-Dtruth|/max %.3g  oracle-vs-truth %.3g  %s %.1f ms" % (
        n, d, nq, k, same, same_t, rel, np.abs(Dr - Dt).max() / Dt.max(), idx.stats(), (t1 - t0) * 1e3), flush=True)
