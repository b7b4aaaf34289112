"""CPU oracle for the IndexFlatL2 path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

PARITY UNPINNED (see flat_l2_oracle.c): faiss-cpu is absent and the reference holds no
golden vectors, so this restates faiss's published behaviour for the calls in
/root/reference/llm/rag.py:61-86,114-120 and is cross-checked against an independent
float64 brute force (`truth_f64`) plus hand-derived known answers (tests/golden/).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this.
"""
import ctypes
import os
import struct

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

FLT_MAX = np.float32(3.4028234663852886e38)


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            import importlib.util
            spec = importlib.util.spec_from_file_location("_oracle_build", os.path.join(_HERE, "build.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build()
        L = ctypes.CDLL(path)
        fp = ctypes.POINTER(ctypes.c_float)
        ip = ctypes.POINTER(ctypes.c_int64)
        L.orc_search.argtypes = [fp, ctypes.c_int64, ctypes.c_int, fp, ctypes.c_int64, ctypes.c_int64, fp, ip, ctypes.c_int]
        L.orc_search.restype = ctypes.c_int
        L.orc_merge.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_int64, fp, ip, fp, ip]
        L.orc_merge.restype = ctypes.c_int
        L.orc_synth.argtypes = [ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, fp]
        L.orc_synth.restype = None
        L.orc_num_threads.restype = ctypes.c_int
        L.orc_set_num_threads.argtypes = [ctypes.c_int]
        L.orc_set_num_threads.restype = None
        L.orc_blas_begin.argtypes = [ctypes.c_int64, ctypes.c_int64]
        L.orc_blas_begin.restype = ctypes.c_void_p
        L.orc_norms.argtypes = [fp, ctypes.c_int64, ctypes.c_int, fp]
        L.orc_norms.restype = None
        L.orc_blas_update.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, fp, fp, fp]
        L.orc_blas_update.restype = None
        L.orc_blas_end.argtypes = [ctypes.c_void_p, fp, ip]
        L.orc_blas_end.restype = None
        _lib = L
    return _lib


def _f(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _i(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


MODE_FAISS, MODE_DIRECT, MODE_EXPAND, MODE_DIRECT_ROWPAR = 0, 1, 2, 3


def search(x, q, k, mode=MODE_FAISS):
    """faiss `IndexFlatL2.search` restated (rag.py:116): returns (D float32 [nq,k], I int64 [nq,k])."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    assert x.ndim == 2 and q.ndim == 2 and (x.shape[0] == 0 or x.shape[1] == q.shape[1])
    assert k > 0
    nq, d = q.shape
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    rc = lib().orc_search(_f(x), x.shape[0], d, _f(q), nq, k, _f(D), _i(I), mode)
    assert rc == 0
    return D, I


def search_blas(x, q, k, db_bs=8192):
    """Rule (2) with the inner-product blocks delegated to the best BLAS present (MKL sgemm via
    torch.mm on the CPU), as faiss delegates them to sgemm_; norms, clamp and result handling in
    the C oracle.  Used as the timed CPU baseline for batched searches; same results as
    search(..., MODE_EXPAND) up to the summation order inside the BLAS."""
    import torch
    x = np.ascontiguousarray(x, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    nq, d = q.shape
    n = x.shape[0]
    L = lib()
    qn = np.empty(nq, np.float32)
    xn = np.empty(max(n, 1), np.float32)
    L.orc_norms(_f(q), nq, d, _f(qn))
    if n:
        L.orc_norms(_f(x), n, d, _f(xn))
    st = L.orc_blas_begin(nq, k)
    qt = torch.from_numpy(q)
    xt = torch.from_numpy(x)
    for j0 in range(0, n, db_bs):
        nb = min(db_bs, n - j0)
        ip = torch.mm(qt, xt[j0:j0 + nb].T).contiguous().numpy()
        L.orc_blas_update(st, nb, j0, _f(ip), _f(qn), _f(xn[j0:j0 + nb]))
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    L.orc_blas_end(st, _f(D), _i(I))
    return D, I


def merge(Dl, Il):
    """Combine per-shard results [nlists,nq,k] into the global top-k (row-sharded search)."""
    Dl = np.ascontiguousarray(Dl, dtype=np.float32)
    Il = np.ascontiguousarray(Il, dtype=np.int64)
    nl, nq, k = Dl.shape
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    lib().orc_merge(nl, nq, k, _f(Dl), _i(Il), _f(D), _i(I))
    return D, I


def synth(seed, row0, nrows, d):
    """Counter-based synthetic rows, bit-identical to the CUDA generator (b2k_synth)."""
    out = np.empty((nrows, d), np.float32)
    if nrows:
        lib().orc_synth(seed, row0, nrows, d, _f(out))
    return out


def num_threads():
    return lib().orc_num_threads()


def use_all_cores():
    """Undo a launcher's OMP_NUM_THREADS=1 (torchrun exports it) for the oracle's OpenMP loops and
    for the BLAS behind `search_blas`; returns the thread count now in use."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(n)
    try:
        import torch
        torch.set_num_threads(n)
    except Exception:
        pass
    return num_threads()


def truth_f64(x, q, k, chunk=262144):
    """Independent ground truth: float64 brute force, (distance, id) lexicographic top-k.

    Returns (D float64 [nq,k] padded with +inf, I int64 padded with -1, Dall-less)."""
    x = np.asarray(x, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    nq = q.shape[0]
    n = x.shape[0]
    bestD = np.full((nq, 0), np.inf)
    bestI = np.full((nq, 0), -1, np.int64)
    qn = (q * q).sum(1)[:, None]
    for j0 in range(0, n, chunk):
        xb = x[j0:j0 + chunk]
        dist = qn + (xb * xb).sum(1)[None, :] - 2.0 * (q @ xb.T)
        # the expansion in float64 keeps ~1e-13 relative accuracy at these magnitudes
        np.maximum(dist, 0.0, out=dist)
        ids = np.arange(j0, j0 + xb.shape[0], dtype=np.int64)[None, :].repeat(nq, 0)
        D = np.concatenate([bestD, dist], 1)
        I = np.concatenate([bestI, ids], 1)
        order = np.lexsort((I, D), axis=1)[:, :k]
        bestD = np.take_along_axis(D, order, 1)
        bestI = np.take_along_axis(I, order, 1)
    if bestD.shape[1] < k:
        pad = k - bestD.shape[1]
        bestD = np.concatenate([bestD, np.full((nq, pad), np.inf)], 1)
        bestI = np.concatenate([bestI, np.full((nq, pad), -1, np.int64)], 1)
    return bestD, bestI


def dist_f64(x, q, ids):
    """float64 squared distances from each query to the rows named in ids [nq,k] (-1 -> inf)."""
    x = np.asarray(x, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    out = np.full(ids.shape, np.inf)
    for i in range(ids.shape[0]):
        m = ids[i] >= 0
        diff = x[ids[i][m]] - q[i][None, :]
        out[i][m] = (diff * diff).sum(1)
    return out


def stream_search(rows, n, d, q, k, chunk=1_000_000, mode=MODE_FAISS):
    """`search` over a row set too large to hold at once: `rows(j0, nb)` yields rows [j0, j0+nb) as
    float32 (e.g. `lambda j0, nb: synth(seed, j0, nb, d)`), each chunk is searched with the rule the
    WHOLE query batch selects (so chunking does not change the numerics: every distance depends on
    its own row only) and the per-chunk results are combined with `merge`.  Same answer as one call
    on all n rows."""
    q = np.ascontiguousarray(q, dtype=np.float32)
    if mode == MODE_FAISS:
        mode = MODE_EXPAND if q.shape[0] >= 20 else MODE_DIRECT
    best = None
    for j0 in range(0, n, chunk):
        nb = min(chunk, n - j0)
        D, I = search(rows(j0, nb), q, k, mode)
        I = np.where(I >= 0, I + j0, I)
        best = (D, I) if best is None else merge(np.stack([best[0], D]), np.stack([best[1], I]))
    if best is None:
        best = search(np.zeros((0, d), np.float32), q, k, mode)
    return best


def compare_knn(D, I, Dref, Iref, row_of, q, rtol=1e-5, expanded=False):
    """Non-raising form of the parity rule (tests/util.py: assert_knn_matches) for result sets over a
    row set that is not in host memory: `row_of(ids)` returns the float32 rows of the given global
    ids.  Distances: |D - Dref| <= rtol * scale (scale = |Dref|, or max(|Dref|, |q|^2 + |x|^2) where
    the reference used the norm expansion); ids: equal, or the float64 distances of the ids that
    differ agree position by position within twice that tolerance (ties at rounding level).
    Returns {"ok", "max_rel_err", "rows_with_other_ids", "why"}."""
    D = np.asarray(D); I = np.asarray(I)
    q = np.asarray(q, dtype=np.float32)
    out = {"ok": False, "max_rel_err": None, "rows_with_other_ids": 0, "why": ""}
    if D.shape != Dref.shape or I.shape != Iref.shape:
        out["why"] = "shape"; return out
    pad = Iref < 0
    if not np.array_equal(I < 0, pad) or not np.all(D[pad] == FLT_MAX):
        out["why"] = "padding differs"; return out
    ids = np.unique(np.concatenate([I[~pad].ravel(), Iref[~pad].ravel()]))
    xs = row_of(ids).astype(np.float64) if ids.size else np.zeros((0, q.shape[1]))
    pos = lambda a: np.searchsorted(ids, a)
    qn = (q.astype(np.float64) ** 2).sum(1)
    xn = (xs * xs).sum(1)
    scale = np.abs(Dref.astype(np.float64))
    if expanded and ids.size:
        scale = np.maximum(scale, qn[:, None] + xn[pos(np.where(pad, ids[0], Iref))])
    err = np.abs(D.astype(np.float64) - Dref.astype(np.float64))
    ok = ~pad
    rel = err[ok] / np.maximum(scale[ok], 1e-30)
    out["max_rel_err"] = float(rel.max()) if rel.size else 0.0
    if rel.size and rel.max() > rtol:
        out["why"] = "distance beyond tolerance"; return out
    if np.any(np.diff(np.where(pad, np.inf, D.astype(np.float64)), axis=1) < 0):
        out["why"] = "D not ascending"; return out
    for i in np.where((I != Iref).any(1))[0]:
        out["rows_with_other_ids"] += 1
        a, b = I[i][~pad[i]], Iref[i][~pad[i]]
        if len(set(a.tolist())) != len(a):
            out["why"] = "duplicate ids in row %d" % i; return out
        da = ((xs[pos(a)] - q[i].astype(np.float64)) ** 2).sum(1)
        db = ((xs[pos(b)] - q[i].astype(np.float64)) ** 2).sum(1)
        t = rtol * np.maximum(np.maximum(da, db), (qn[i] + xn[pos(a)]) if expanded else 0) + 1e-30
        if np.any(np.abs(da - db) > 2 * t):
            out["why"] = "row %d: ids differ beyond the tie tolerance" % i; return out
    out["ok"] = True
    return out


def synth_rows(seed, d):
    """row_of callable for compare_knn over the synthetic stream."""
    def row_of(ids):
        out = np.empty((len(ids), d), np.float32)
        for j, g in enumerate(ids):
            out[j] = synth(seed, int(g), 1, d)[0]
        return out
    return row_of


# --------------------------------------------------------------------------------------
# IxF2: the file faiss.write_index / read_index produce for an IndexFlatL2
# (rag.py:62,74,82,153,186).  Little-endian, no padding [faiss-upstream, SURVEY.md 8a-a5]:
#   "IxF2" | d:int32 | ntotal:int64 | 1<<20:int64 | 1<<20:int64 | is_trained:uint8 |
#   metric_type:int32 (1 = L2) | count:uint64 (= ntotal*d) | count float32 values

_HDR = struct.Struct("<4siqqqBiQ")
assert _HDR.size == 45


def write_ixf2(path, x, d=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    if d is None:
        d = x.shape[1]
    n = x.shape[0] if x.size else 0
    with open(path, "wb") as f:
        f.write(_HDR.pack(b"IxF2", d, n, 1 << 20, 1 << 20, 1, 1, n * d))
        if n:
            f.write(x.tobytes())


def read_ixf2(path):
    with open(path, "rb") as f:
        hdr = f.read(45)
        if len(hdr) != 45:
            raise RuntimeError("short IxF2 header")
        four, d, n, _d0, _d1, trained, metric, count = _HDR.unpack(hdr)
        if four != b"IxF2":
            raise RuntimeError("not an IxF2 file: %r" % four)
        if metric != 1 or count != n * d:
            raise RuntimeError("inconsistent IxF2 header")
        data = np.frombuffer(f.read(4 * count), dtype=np.float32)
        if data.size != count:
            raise RuntimeError("truncated IxF2 payload")
    return d, data.reshape(n, d).copy()


def empty_ixf2_bytes(d):
    """The 45 bytes SURVEY.md 8c lists for an empty index of dimension d."""
    return (b"IxF2" + struct.pack("<i", d) + b"\x00" * 8 + b"\x00\x00\x10\x00\x00\x00\x00\x00" * 2
            + b"\x01" + b"\x01\x00\x00\x00" + b"\x00" * 8)
