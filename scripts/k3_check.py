"""Quick GPU check of the tensor-core path against the oracle (run under `timeout`).
usage: k3_check.py [cg=1|2] n,d,nq,k ..."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import b200knn
from oracle import oracle as O
from util import unit_rows

args = sys.argv[1:]
cg = 1
unit = False
sep = 0
tmode = 0
seg = 1
ew = 0
while args and "=" in args[0]:
    kx, v = args.pop(0).split("=")
    if kx == "cg": cg = int(v)
    if kx == "unit": unit = bool(int(v))
    if kx == "sep": sep = int(v)
    if kx == "mode": tmode = int(v)
    if kx == "seg": seg = int(v)
    if kx == "ew": ew = int(v)
cases = [(1000, 64, 128, 10), (5000, 384, 40, 10), (100000, 384, 300, 100), (70000, 100, 1024, 5), (3001, 1024, 129, 37)]
if args:
    cases = [tuple(int(v) for v in a.split(",")) for a in args]
for (n, d, nq, k) in cases:
    if n <= 2000000:
        x = O.synth(1234, 0, n, d); q = O.synth(5678, 0, nq, d)
        if unit: x, q = unit_rows(x), unit_rows(q)
        idx = b200knn.IndexFlatL2(d); idx.add(x)
    else:
        x = None; q = b200knn.synth(5678, 0, nq, d).cpu().numpy()
        idx = b200knn.IndexFlatL2(d); idx.reserve(n); idx.add_synthetic(n, 1234, 0)
    idx.set_option("path", 2); idx.set_option("timing", 1); idx.set_option("tensor_cta_group", cg); idx.set_option("tensor_sep_accum", sep); idx.set_option("tensor_mode", tmode); idx.set_option("tensor_segments", seg); idx.set_option("tensor_epilogue_warps", ew)
    D, I = idx.search(q, k)
    t0 = time.time(); D, I = idx.search(q, k); t1 = time.time()
    st = idx.stats()
    tf = 6.0 * nq * n * d / (st["main_kernel_ms"] * 1e-3) / 1e12
    if x is not None and n * nq <= 4e8:
        Dr, Ir = O.search(x, q, k, O.MODE_EXPAND)
        Dt, It = O.truth_f64(x, q, k)
        qn = (q.astype(np.float64) ** 2).sum(1)[:, None]; xn = (x.astype(np.float64) ** 2).sum(1)
        scale = np.maximum(Dt, qn + xn[np.maximum(It, 0)])
        ok = It >= 0
        rel = (np.abs(D - Dt)[ok] / scale[ok]).max()
        relo = (np.abs(Dr - Dt)[ok] / scale[ok]).max()
        print("mode=%d sep=%d " % (tmode, sep), end=""); print("cg=%d n=%d d=%d nq=%d k=%d  ids==oracle %.5f ids==f64 %.5f  max rel err (vs cancel scale) %.3g (oracle %.3g)  kernel %.3f ms = %.0f TF/s, call %.1f ms" % (
            cg, n, d, nq, k, (I == Ir).mean(), (I == It).mean(), rel, relo, st["main_kernel_ms"], tf, (t1 - t0) * 1e3), flush=True)
    else:
        print("mode=%d cg=%d n=%d d=%d nq=%d k=%d  kernel %.3f ms, call %.1f ms" % (tmode, cg, n, d, nq, k, st["main_kernel_ms"], (t1 - t0) * 1e3), flush=True)
