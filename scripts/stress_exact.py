"""Exactness stress for the screening paths (bf16 / TF32 tensor screening, bf16 scan): distributions that
are hard for a reduced-precision screen, checked against the float64 ground truth.
    python scripts/stress_exact.py
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import b200knn
from oracle import oracle as O


def cases(rng):
    n, d = 120_000, 128
    g = rng.standard_normal((n, d)).astype(np.float32)
    yield "gaussian", g, rng.standard_normal((64, d)).astype(np.float32)
    s = np.exp(rng.standard_normal((n, 1)) * 1.5).astype(np.float32)
    yield "heavy-tailed row norms", g * s, rng.standard_normal((64, d)).astype(np.float32)
    yield "offset mean (100 + N(0,1))", (g + 100.0).astype(np.float32), (rng.standard_normal((64, d)) + 100.0).astype(np.float32)
    c = rng.standard_normal((300, d)).astype(np.float32)
    cl = (c[rng.integers(0, 300, n)] + rng.standard_normal((n, d)).astype(np.float32) * 1e-3).astype(np.float32)
    yield "tight clusters (sigma 1e-3)", cl, (c[:64] + rng.standard_normal((64, d)).astype(np.float32) * 1e-3).astype(np.float32)
    u = g / np.linalg.norm(g, axis=1, keepdims=True)
    qs = u[rng.integers(0, n, 64)] + rng.standard_normal((64, d)).astype(np.float32) * 0.01
    yield "unit norm, queries near stored rows", u.astype(np.float32), (qs / np.linalg.norm(qs, axis=1, keepdims=True)).astype(np.float32)
    tiny = (g * 1e-18).astype(np.float32)
    yield "tiny magnitudes (1e-18)", tiny, (rng.standard_normal((64, d)) * 1e-18).astype(np.float32)
    big = (g * 1e15).astype(np.float32)
    yield "huge magnitudes (1e15)", big, (rng.standard_normal((64, d)) * 1e15).astype(np.float32)
    sp = g * (rng.random((n, d)) < 0.05)
    yield "sparse rows (5 % non-zero)", sp.astype(np.float32), (rng.standard_normal((64, d)) * (rng.random((64, d)) < 0.05)).astype(np.float32)
    lat = rng.integers(-3, 4, (n, d)).astype(np.float32)
    yield "integer lattice (many exact ties)", lat, rng.integers(-3, 4, (64, d)).astype(np.float32)


def check(name, what, D, I, x, q, k):
    Dt, It = O.truth_f64(x, q, k)
    dd = O.dist_f64(x, q, I)
    # every returned row must be as close as the true k-th neighbour up to fp32 rounding of the distance
    kth = Dt[:, k - 1:k]
    tol = 4e-6 * np.maximum(kth, 1e-300) + 1e-37
    ok_rows = np.all(dd <= kth + tol, axis=1)
    # and the reported distances are the fp32 direct sums of those rows
    rel = np.abs(D.astype(np.float64) - dd) / np.maximum(dd, 1e-300)
    rel = np.where(dd > 1e-37, rel, 0.0)
    ok_d = rel.max() < 4e-6
    uniq = all(len(set(r.tolist())) == k for r in I)
    same = np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(I, It)])
    print("%-38s %-22s rows ok %s  dist ok %s (max rel %.2g)  unique %s  overlap with f64 top-k %.5f" % (name, what, ok_rows.all(), ok_d, rel.max(), uniq, same), flush=True)
    return ok_rows.all() and ok_d and uniq


rng = np.random.default_rng(7)
bad = 0
for name, x, q in cases(rng):
    idx = b200knn.IndexFlatL2(x.shape[1]); idx.add(x)
    for what, opts, qq, k in (("tensor bf16 screen", {"path": 2, "tensor_mode": 3}, q, 20),
                              ("tensor tf32 screen", {"path": 2, "tensor_mode": 2}, q, 20),
                              ("scan bf16 + rerank", {"path": 1, "scan_rows": 2}, q[:3], 20),
                              ("scan fp32", {"path": 1, "scan_rows": 1}, q[:3], 20)):
        for kk, v in opts.items(): idx.set_option(kk, v)
        D, I = idx.search(qq, k)
        st = idx.stats()
        extra = " [%s%s]" % (st.get("tensor_mode") or st.get("scan_rows"), ", fallbacks %s" % st.get("screen_fallbacks") if st.get("screen_fallbacks") else "")
        if not check(name, what + extra, D, I, x, qq, k): bad += 1
print("FAILED %d" % bad if bad else "ALL OK")
