import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, b200knn
from oracle import oracle as O
rng = np.random.default_rng(11)
n, d, nq, k = 700_000, 64, 512, 10
x = rng.standard_normal((n, d)).astype(np.float32)
big = rng.choice(n, 20, replace=False)
x[big] *= 30.0
q = rng.standard_normal((nq, d)).astype(np.float32)
q[0] = x[big[0]]; q[1] = x[big[1]] + 0.5 * rng.standard_normal(d).astype(np.float32); q[2] = 25.0 * rng.standard_normal(d).astype(np.float32)
idx = b200knn.IndexFlatL2(d); idx.add(x)
D, I = idx.search(q, k); print(idx.stats())
y = rng.standard_normal((5000, d)).astype(np.float32)
y[::500] *= 40.0
idx.add(y)
xa = np.concatenate([x, y])
q[3] = y[500]
for trial in range(2):
    D, I = idx.search(q, k); print(idx.stats())
    Dt, It = O.truth_f64(xa, q, k)
    d64 = O.dist_f64(xa, q, I)
    bad = np.argwhere(np.abs(d64 - Dt) > 1e-5 * np.maximum(Dt, 1e-30) + 1e-30)
    print("bad entries", len(bad), bad[:10].tolist())
    norms = (xa.astype(np.float64) ** 2).sum(1)
    for (qi, j) in bad[:6]:
        missing = sorted(set(It[qi].tolist()) - set(I[qi].tolist()))
        print(" q", qi, "rank", j, "got", I[qi].tolist(), "want", It[qi].tolist(), "missing", missing, "norm2 of missing", [round(norms[m], 1) for m in missing], "|q|2", round(float((q[qi].astype(np.float64) ** 2).sum()), 1))
