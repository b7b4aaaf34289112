"""GPU parity tests: the CUDA path, called through the C ABI (via the ctypes shim), against the
CPU oracle on identical inputs.  The reference has no tests; the cases
follow SURVEY.md section 7 step 2 / section 8c and the call pattern of llm/rag.py.

Tolerance (north_star): distances within 1e-5 relative; ids identical except where the
competing distances agree within that tolerance.  Integer-lattice cases are bit-exact."""
import os
import pickle

import numpy as np
import pytest

from oracle import oracle as O
from util import RTOL, assert_knn_matches, unit_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def faiss():
    import b200knn
    return b200knn


def _build(faiss, x):
    idx = faiss.IndexFlatL2(x.shape[1])
    if x.shape[0]:
        idx.add(x)
    return idx


def test_synth_generator_is_bit_identical_to_oracle(faiss):
    for (seed, row0, n, d) in [(42, 7, 4, 8), (1234, 0, 1000, 384), (5678, 123456789, 33, 1027)]:
        g = faiss.synth(seed, row0, n, d).cpu().numpy()
        assert np.array_equal(g, O.synth(seed, row0, n, d))


def test_config1_rag_default_2048x384_nq1_k5(faiss):
    """BASELINE.json configs[0]: the reference's own CPU-runnable case."""
    x = O.synth(1234, 0, 2048, 384)
    q = O.synth(5678, 0, 1, 384)
    idx = _build(faiss, x)
    assert idx.ntotal == 2048 and idx.d == 384
    D, I = idx.search(q, 5)
    assert_knn_matches(D, I, x, q, 5, what="C1")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "c1_2048x384_nq1_k5.npz"))
    assert np.array_equal(I, g["I"]) and np.allclose(D, g["D"], rtol=1e-5)


def test_real_app_shape_unit_norm_64x1024(faiss):
    """What rag.py really feeds: bge-m3 d=1024, normalize_embeddings=True (rag.py:33,55), tens of rows."""
    x = unit_rows(O.synth(1234, 0, 64, 1024))
    q = unit_rows(O.synth(5678, 0, 1, 1024))
    D, I = _build(faiss, x).search(q, 5)
    assert_knn_matches(D, I, x, q, 5, what="app")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "app_64x1024_unit_nq1_k5.npz"))
    assert np.array_equal(I, g["I"]) and np.allclose(D, g["D"], rtol=1e-5)


@pytest.mark.parametrize("d", [1, 3, 4, 100, 384, 768, 1024, 1027])
def test_dimensions(faiss, d):
    rng = np.random.default_rng(d)
    x = rng.standard_normal((3001, d)).astype(np.float32)
    q = rng.standard_normal((3, d)).astype(np.float32)
    D, I = _build(faiss, x).search(q, 10)
    assert_knn_matches(D, I, x, q, 10, what="d=%d" % d)


@pytest.mark.parametrize("k", [1, 5, 10, 100, 129, 1024])
def test_k_values(faiss, k):
    rng = np.random.default_rng(k)
    x = rng.standard_normal((5000, 64)).astype(np.float32)
    q = rng.standard_normal((2, 64)).astype(np.float32)
    D, I = _build(faiss, x).search(q, k)
    assert_knn_matches(D, I, x, q, k, what="k=%d" % k)


@pytest.mark.parametrize("nq", [1, 2, 7, 8, 9, 19])
def test_small_query_batches_use_direct_numerics(faiss, nq):
    rng = np.random.default_rng(nq)
    x = rng.standard_normal((20000, 96)).astype(np.float32)
    q = rng.standard_normal((nq, 96)).astype(np.float32)
    D, I = _build(faiss, x).search(q, 10)
    assert_knn_matches(D, I, x, q, 10, what="nq=%d" % nq)


@pytest.mark.parametrize("nq", [20, 33, 130, 300])
def test_query_batches(faiss, nq):
    rng = np.random.default_rng(nq)
    x = rng.standard_normal((20000, 128)).astype(np.float32)
    q = rng.standard_normal((nq, 128)).astype(np.float32)
    D, I = _build(faiss, x).search(q, 10)
    assert_knn_matches(D, I, x, q, 10, expanded=True, what="nq=%d" % nq)


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 127, 128, 129])
def test_fewer_rows_than_k_pads_with_minus_one(faiss, n):
    """rag.py starts from an empty index (rag.py:61) and searches with k=5 (rag.py:114): the
    -1 / FLT_MAX padding must be exact because rag.py:119 indexes metadata with it."""
    x = O.synth(1, 0, n, 32)
    q = O.synth(2, 0, 2, 32)
    idx = _build(faiss, x)
    assert idx.ntotal == n
    D, I = idx.search(q, 5)
    Dr, Ir = O.search(x, q, 5)
    assert np.array_equal(I, Ir)
    assert np.array_equal(D == O.FLT_MAX, Dr == O.FLT_MAX)
    assert_knn_matches(D, I, x, q, 5, Dr, Ir, what="n=%d" % n)


def test_integer_lattice_is_bit_exact(faiss):
    rng = np.random.default_rng(7)
    x = rng.integers(-8, 9, size=(30000, 48)).astype(np.float32)
    for nq in (1, 5, 25, 140):
        q = rng.integers(-8, 9, size=(nq, 48)).astype(np.float32)
        D, I = _build(faiss, x).search(q, 20)
        Dr, Ir = O.search(x, q, 20)
        assert np.array_equal(D, Dr) and np.array_equal(I, Ir), nq


def test_all_identical_rows_tie_break_by_id(faiss):
    x = np.ones((5000, 8), np.float32)
    for nq in (1, 24):
        q = np.zeros((nq, 8), np.float32)
        D, I = _build(faiss, x).search(q, 7)
        assert np.all(I == np.arange(7)[None, :]) and np.all(D == 8.0)


def test_duplicates_and_self_match(faiss):
    rng = np.random.default_rng(5)
    base = unit_rows(rng.standard_normal((500, 1024)))
    x = np.concatenate([base, base[:100]], 0)          # rows 500..599 duplicate rows 0..99
    q = base[[3, 42]].copy()
    D, I = _build(faiss, x).search(q, 4)
    assert I[:, :2].tolist() == [[3, 503], [42, 542]]
    assert np.all(D[:, :2] == 0.0)                     # direct differences: exact zero
    assert_knn_matches(D, I, x, q, 4, what="dups")
    qb = np.repeat(q, 12, 0)                           # batched numerics: >= 0, never negative
    D, I = _build(faiss, x).search(qb, 4)
    assert np.all(D >= 0.0)
    assert_knn_matches(D, I, x, qb, 4, expanded=True, what="dups-batched")


def test_nonfinite_rows_are_never_selected(faiss):
    x = np.array([[1.0, 0, 0, 0], [np.inf, 0, 0, 0], [np.nan, 0, 0, 0], [2.0, 0, 0, 0]], np.float32)
    D, I = _build(faiss, x).search(np.zeros((1, 4), np.float32), 4)
    assert I.tolist() == [[0, 3, -1, -1]]


def test_incremental_adds_keep_sequential_ids(faiss):
    x = O.synth(77, 0, 5000, 64)
    idx = faiss.IndexFlatL2(64)
    for a, b in ((0, 1), (1, 1500), (1500, 1501), (1501, 5000)):   # forces several regrowths
        idx.add(x[a:b])
    assert idx.ntotal == 5000
    q = O.synth(78, 0, 4, 64)
    D, I = idx.search(q, 10)
    assert_knn_matches(D, I, x, q, 10, what="incremental")
    assert np.array_equal(idx.reconstruct_n(0, 5000), x)
    idx.reset()
    assert idx.ntotal == 0 and idx.d == 64
    D, I = idx.search(q, 3)
    assert np.all(I == -1)
    idx.add(x[:10])
    assert idx.search(x[:1], 1)[1].tolist() == [[0]]


def test_shape_and_k_errors_are_assertions(faiss):
    idx = faiss.IndexFlatL2(16)
    with pytest.raises(AssertionError):
        idx.add(np.zeros((3, 8), np.float32))
    with pytest.raises(AssertionError):
        idx.search(np.zeros((1, 8), np.float32), 5)
    with pytest.raises(AssertionError):
        idx.search(np.zeros((1, 16), np.float32), 0)
    with pytest.raises(RuntimeError):
        idx.search(np.zeros((1, 16), np.float32), 1025)
    idx.add(np.zeros((2, 16), np.float64))               # coerced to float32 like faiss


def test_ixf2_bytes_and_round_trip(faiss, tmp_path):
    p = str(tmp_path / "faiss_index")
    faiss.write_index(faiss.IndexFlatL2(1024), p)
    assert open(p, "rb").read() == O.empty_ixf2_bytes(1024)
    back = faiss.read_index(p)
    assert back.ntotal == 0 and back.d == 1024
    x = O.synth(5, 0, 1234, 100)
    idx = _build(faiss, x)
    faiss.write_index(idx, p)
    ref = str(tmp_path / "ref")
    O.write_ixf2(ref, x)
    assert open(p, "rb").read() == open(ref, "rb").read()  # byte-identical to the independent writer
    assert not os.path.exists(p + ".tmp")
    back = faiss.read_index(ref)
    assert back.ntotal == 1234 and back.d == 100
    q = O.synth(6, 0, 3, 100)
    D, I = back.search(q, 5)
    assert_knn_matches(D, I, x, q, 5, what="after read_index")


def test_rag_py_call_pattern(faiss, tmp_path):
    rag_py_call_pattern(faiss, tmp_path)


def rag_py_call_pattern(faiss, tmp_path):
    """ensure_index_exists / update_index / search_documents restated (rag.py:57-86,114-120) with
    the encoder stubbed: row id == index into the metadata list, -1 padding reaches metadata[-1]."""
    INDEX_PATH = str(tmp_path / "faiss_index")
    d = 1024
    enc = lambda i: unit_rows(O.synth(900, i, 1, d))[0]

    def ensure_index_exists():
        if not os.path.exists(INDEX_PATH):
            faiss.write_index(faiss.IndexFlatL2(d), INDEX_PATH)
            pickle.dump([], open(INDEX_PATH + ".metadata", "wb"))

    def update_index(vectors, metadata):
        index = faiss.read_index(INDEX_PATH)
        existing = pickle.load(open(INDEX_PATH + ".metadata", "rb"))
        vectors_np = np.array(vectors).astype("float32")
        if vectors_np.shape[0] > 0:
            index.add(vectors_np)
            existing.extend(metadata)
            faiss.write_index(index, INDEX_PATH)
            pickle.dump(existing, open(INDEX_PATH + ".metadata", "wb"))
        return index.ntotal, index.d

    def search_documents(qv, k=5):
        index = faiss.read_index(INDEX_PATH)
        metadata = pickle.load(open(INDEX_PATH + ".metadata", "rb"))
        D, I = index.search(qv.reshape(1, -1), k)
        return [(metadata[idx], float(dist)) for idx, dist in zip(I[0], D[0])]

    ensure_index_exists()
    with pytest.raises(IndexError):                       # ntotal == 0 -> metadata[-1] on [] (rag.py:119)
        search_documents(enc(0))
    assert update_index([enc(i) for i in range(3)], [{"chunk_id": i} for i in range(3)]) == (3, d)
    res = search_documents(enc(1))
    assert res[0] == ({"chunk_id": 1}, 0.0)
    assert res[3][0] == {"chunk_id": 2} and res[3][1] == float(O.FLT_MAX)   # idx -1 -> last chunk
    assert update_index([enc(i) for i in range(3, 40)], [{"chunk_id": i} for i in range(3, 40)]) == (40, d)
    res = search_documents(enc(17))
    assert res[0][0] == {"chunk_id": 17} and res[0][1] == 0.0 and len(res) == 5
    assert isinstance(faiss.read_index(INDEX_PATH).ntotal, int)


def test_torch_device_tensors(faiss):
    import torch
    x = faiss.synth(1234, 0, 50000, 384)
    q = faiss.synth(5678, 0, 4, 384)
    idx = faiss.IndexFlatL2(384)
    idx.add(x)
    D, I = idx.search(q, 10)
    assert D.is_cuda and I.dtype == torch.int64
    xh, qh = x.cpu().numpy(), q.cpu().numpy()
    assert_knn_matches(D.cpu().numpy(), I.cpu().numpy(), xh, qh, 10, what="torch")
    idx2 = faiss.IndexFlatL2(384)
    idx2.add_synthetic(50000, 1234, 0)
    D2, I2 = idx2.search(qh, 10)
    assert np.array_equal(I2, I.cpu().numpy()) and np.array_equal(D2, D.cpu().numpy())


def test_merge_topk_matches_oracle(faiss):
    rng = np.random.default_rng(2)
    nl, nq, k = 8, 5, 100
    Dl = np.sort(rng.random((nl, nq, k)).astype(np.float32), axis=2)
    Dl[:, :, ::7] = Dl[:, :, 1::7][:, :, :Dl[:, :, ::7].shape[2]]      # inject exact ties
    Dl = np.sort(Dl, axis=2)
    Il = rng.permutation(nl * nq * k).reshape(nl, nq, k).astype(np.int64)
    Il[3, :, 90:] = -1
    Dl[3, :, 90:] = O.FLT_MAX
    Dm, Im = faiss.merge_topk(Dl, Il)
    Dr, Ir = O.merge(Dl, Il)
    assert np.array_equal(Dm, Dr) and np.array_equal(Im, Ir)
    off = np.arange(nl, dtype=np.int64) * 1000000
    Dm, Im = faiss.merge_topk(Dl, Il, id_offset=off)
    Dr, Ir = O.merge(Dl, np.where(Il >= 0, Il + off[:, None, None], -1))
    assert np.array_equal(Dm, Dr) and np.array_equal(Im, Ir)


def test_one_million_rows_nq1_k10(faiss):
    """BASELINE.json configs[1] at full size: 1M x 384, nq=1, k=10, against the oracle (about a
    second of CPU) and by a size-independent property: searching the two halves separately and
    merging gives the same answer."""
    idx = faiss.IndexFlatL2(384)
    idx.add_synthetic(1_000_000, 1234, 0)
    x = O.synth(1234, 0, 1_000_000, 384)
    q = O.synth(5678, 0, 1, 384)
    D, I = idx.search(q, 10)
    assert_knn_matches(D, I, x, q, 10, what="C2")
    halves = []
    for a, b in ((0, 500_000), (500_000, 1_000_000)):
        h = faiss.IndexFlatL2(384)
        h.add(x[a:b])
        halves.append(h.search(q, 10))
    Dm, Im = faiss.merge_topk(np.stack([h[0] for h in halves]), np.stack([h[1] for h in halves]),
                              id_offset=np.array([0, 500_000]))
    assert np.array_equal(Im, I) and np.array_equal(Dm, D)


@pytest.mark.parametrize("mode,cg,sep,k", [(1, 1, 2, 10), (1, 1, 1, 10), (1, 2, 2, 100), (1, 2, 1, 100), (1, 2, 2, 700),
                                           (1, 2, 1, 1024), (2, 1, 0, 10), (2, 1, 0, 700), (2, 2, 0, 100), (2, 2, 0, 1024),
                                           (3, 1, 0, 10), (3, 1, 0, 100), (3, 1, 0, 700), (3, 1, 0, 1024),
                                           (3, 2, 0, 100), (4, 1, 0, 10), (4, 2, 0, 100), (4, 2, 0, 700)])
def test_tensor_kernel_variants(faiss, mode, cg, sep, k):
    """Every instantiation of the tcgen05 kernel -- 3xTF32 (one CTA / CTA pair, shared / separate
    cross-term accumulator, short / long lists) and TF32 / bf16 / 3xTF32 screening + exact re-rank -- against the
    oracle's expansion numerics, on N(0,1) rows and on unit-norm rows with planted duplicates (the
    rag.py distribution)."""
    rng = np.random.default_rng(cg * 10 + sep)
    for unit in (False, True):
        x = O.synth(31, 0, 40000, 256)
        q = O.synth(32, 0, 70, 256)
        if unit:
            x, q = unit_rows(x), unit_rows(q)
            q[:8] = x[rng.integers(0, 40000, 8)]          # self matches: true distance 0
        idx = _build(faiss, x)
        idx.set_option("path", 2)
        idx.set_option("tensor_mode", mode)
        idx.set_option("tensor_cta_group", cg)
        idx.set_option("tensor_sep_accum", sep)
        D, I = idx.search(q, k)
        st = idx.stats()
        assert st["path"] == "tensor" and st["tensor_mode"] == {1: "3xtf32", 2: "screen+rerank", 3: "bf16screen+rerank", 4: "3xtf32screen+rerank"}[mode]
        assert np.all(D >= 0.0)
        if mode >= 2 and unit:
            assert np.all(D[:8, 0] == 0.0)                # exact re-rank: a stored query is at distance exactly 0
        assert_knn_matches(D, I, x, q, k, expanded=True, what="mode=%d cg=%d sep=%d k=%d unit=%s" % (mode, cg, sep, k, unit))


def test_screening_falls_back_when_the_slack_zone_overflows(faiss):
    """Screening keeps every row whose approximate distance is within the error slack of the k-th;
    thousands of rows at (almost) the same distance do not fit its lists, and the search must then
    repeat itself with a narrower screen (3xTF32 + re-rank) and, if that zone is crowded too, with the
    exact scan kernel, instead of dropping candidates."""
    rng = np.random.default_rng(0)
    base = rng.standard_normal((1, 64)).astype(np.float32)
    x = np.repeat(base, 20000, 0) + rng.standard_normal((20000, 64)).astype(np.float32) * 1e-6   # near-duplicates
    x[::1000] = rng.standard_normal((20, 64)).astype(np.float32)
    q = rng.standard_normal((30, 64)).astype(np.float32)
    idx = _build(faiss, x)
    D, I = idx.search(q, 10)
    st = idx.stats()
    # 20000 rows within 1e-6 of each other crowd even the 3xTF32 screen's zone: the scan kernel answers
    assert st["screen_fallbacks"] >= 1 and (st["path"] == "scan" or st.get("tensor_mode") == "3xtf32screen+rerank")
    assert_knn_matches(D, I, x, q, 10, expanded=True, what="overflow fallback")


def test_full_size_c3_properties(faiss):
    """BASELINE.json configs[2] at full size (10M x 384, nq=1024, k=100), checked through
    size-independent properties: (a) rows sorted, ids unique and in range; (b) searching the two
    halves separately and merging gives the same answer; (c) for a few queries the exact scan
    kernel (direct differences, path=1) returns the same neighbours within tolerance; (d) the
    reported distances match float64 distances recomputed on the CPU for the returned ids."""
    import torch
    N, d, nq, k = 10_000_000, 384, 1024, 100
    idx = faiss.IndexFlatL2(d)
    idx.reserve(N)
    idx.add_synthetic(N, 1234, 0)
    q = faiss.synth(5678, 0, nq, d)
    D, I = idx.search(q, k)
    assert idx.stats()["path"] == "tensor"
    Dh, Ih = D.cpu().numpy(), I.cpu().numpy()
    assert np.all(np.diff(Dh, axis=1) >= 0) and Ih.min() >= 0 and Ih.max() < N
    assert all(len(set(r.tolist())) == k for r in Ih[::64])
    # (b) halves + merge
    parts = []
    for a in (0, N // 2):
        h = faiss.IndexFlatL2(d)
        h.reserve(N // 2)
        h.add_synthetic(N // 2, 1234, a)
        parts.append(h.search(q, k))
        del h
    Dm, Im = faiss.merge_topk(torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts]),
                              id_offset=np.array([0, N // 2]))
    assert torch.equal(Im, I) and torch.equal(Dm, D)
    # (c) exact scan on a few queries
    idx.set_option("path", 1)
    sel = [0, 17, 513, 1023]
    Ds, Is = idx.search(q[sel].contiguous(), k)
    Ds, Is = Ds.cpu().numpy(), Is.cpu().numpy()
    qh = q.cpu().numpy()
    for j, qi in enumerate(sel):
        assert np.allclose(Dh[qi], Ds[j], rtol=RTOL)
        differ = Ih[qi] != Is[j]
        if differ.any():       # only near-ties may swap
            assert np.allclose(Dh[qi][differ], Ds[j][differ], rtol=2 * RTOL)
            assert len(set(Ih[qi]) ^ set(Is[j])) <= 2
        # (d) float64 recomputation for the returned ids
        rows = np.stack([O.synth(1234, int(r), 1, d)[0] for r in Ih[qi]])
        d64 = ((rows.astype(np.float64) - qh[qi].astype(np.float64)) ** 2).sum(1)
        mag = (rows.astype(np.float64) ** 2).sum(1) + (qh[qi].astype(np.float64) ** 2).sum()
        assert np.all(np.abs(d64 - Dh[qi]) <= RTOL * np.maximum(d64, mag))


def test_large_k_large_batch_stress(faiss):
    """The shape of BASELINE.json configs[4] (d=1024, nq=4096, k=1024) at 300k rows: against the
    exact scan kernel on a sample of queries, plus the halves-merge property."""
    import torch
    N, d, nq, k = 300_000, 1024, 4096, 1024
    idx = faiss.IndexFlatL2(d)
    idx.add_synthetic(N, 1234, 0)
    q = faiss.synth(5678, 0, nq, d)
    D, I = idx.search(q, k)
    assert idx.stats()["path"] == "tensor"
    idx.set_option("path", 1)
    sel = [0, 1000, 4095]
    Ds, Is = idx.search(q[sel].contiguous(), k)
    for j, qi in enumerate(sel):
        a, b = D[qi].cpu().numpy(), Ds[j].cpu().numpy()
        assert np.allclose(a, b, rtol=RTOL, atol=RTOL * 2048)
        assert len(set(I[qi].cpu().tolist()) ^ set(Is[j].cpu().tolist())) <= 4


def _stat(idx, key):
    import ctypes as C
    from b200knn import _lib
    v = C.c_double()
    _lib.check(_lib.lib().b2k_get_stat(idx._h, key.encode(), C.byref(v)))
    return v.value


def test_resident_index_cache_and_incremental_write(faiss, tmp_path, monkeypatch):
    """rag.py re-reads the index file on every request (rag.py:153) and
    rewrites it on every upload (rag.py:74-84).  An unchanged file must come back resident (its rows
    shared in HBM, no reload) but, as with faiss, as an INDEPENDENT index each time; an index that
    only grew since it was loaded is appended to the file, and the file stays byte-identical to a
    full write; any other change falls back to load / full write."""
    monkeypatch.delenv("B200KNN_INDEX_CACHE", raising=False)
    faiss.clear_index_cache()
    p = str(tmp_path / "faiss_index")
    x = O.synth(3, 0, 5000, 64)
    idx = faiss.IndexFlatL2(64)
    idx.add(x[:2000])
    faiss.write_index(idx, p)
    a = faiss.read_index(p)
    b = faiss.read_index(p)
    assert a is not b and a.ntotal == b.ntotal == 2000
    assert _stat(a, "shares_rows") == 1 and _stat(b, "shares_rows") == 1      # resident: no second load
    a.add(x[2000:3500])                             # update_index: read -> add -> write
    assert a.ntotal == 3500 and b.ntotal == 2000    # ... invisible to the request that holds b
    q = O.synth(4, 0, 2, 64)
    Db, Ib = b.search(q, 3)
    assert_knn_matches(Db, Ib, x[:2000], q, 3, what="other view after add")
    c = faiss.read_index(p)                         # the file still has 2000 rows
    assert c.ntotal == 2000
    b.add(x[4000:4100])                             # a second writer: the shared rows have moved on -> private copy
    assert b.ntotal == 2100 and a.ntotal == 3500 and c.ntotal == 2000
    assert np.array_equal(b.reconstruct_n(1990, 110), np.concatenate([x[1990:2000], x[4000:4100]]))
    assert np.array_equal(a.reconstruct_n(1990, 20), x[1990:2010])
    size_before = os.path.getsize(p)
    faiss.write_index(a, p)                         # appended in place
    assert os.path.getsize(p) == 45 + 4 * 3500 * 64 and os.path.getsize(p) > size_before
    ref = str(tmp_path / "ref")
    O.write_ixf2(ref, x[:3500])
    assert open(p, "rb").read() == open(ref, "rb").read()
    f = faiss.read_index(p)                         # cache follows the write: no load, rows shared with a
    assert f.ntotal == 3500 and _stat(f, "shares_rows") == 1
    a.add(x[3500:3600])                             # ... and what a does next does not reach f
    assert f.ntotal == 3500
    Df, If = f.search(q, 3)
    assert_knn_matches(Df, If, x[:3500], q, 3, what="view of the written file")
    d2, x2 = O.read_ixf2(p)
    assert d2 == 64 and np.array_equal(x2, x[:3500])
    # a file changed behind our back is reloaded
    O.write_ixf2(p, x[:100])
    os.utime(p, ns=(1, 1))
    e = faiss.read_index(p)
    assert e.ntotal == 100
    # a reset index is written in full, not appended
    a.reset()
    a.add(x[:10])
    faiss.write_index(a, p)
    assert os.path.getsize(p) == 45 + 4 * 10 * 64
    D, I = faiss.read_index(p).search(q, 3)
    assert_knn_matches(D, I, x[:10], q, 3, what="after rewrite")
    assert f.ntotal == 3500 and np.array_equal(f.reconstruct_n(3490, 10), x[3490:3500])   # old views live on
    # cache off: independent loads, as faiss
    monkeypatch.setenv("B200KNN_INDEX_CACHE", "0")
    g1, g2 = faiss.read_index(p), faiss.read_index(p)
    assert g1 is not g2 and _stat(g1, "shares_rows") == 0


def test_requests_racing_an_upload_see_consistent_indexes(faiss, tmp_path, monkeypatch):
    """Flask's dev server is threaded (rag.py:204): /generate (rag.py:153-157: read_index + metadata pickle +
    search) runs while /upload_pdf (rag.py:141 -> update_index: read, add, write) runs.  Every request must
    get an index whose rows match SOME state of the file, never rows it has no metadata for."""
    import threading
    monkeypatch.delenv("B200KNN_INDEX_CACHE", raising=False)
    faiss.clear_index_cache()
    p = str(tmp_path / "faiss_index")
    d = 32
    x = O.synth(9, 0, 4000, d)
    idx = faiss.IndexFlatL2(d)
    idx.add(x[:1000])
    faiss.write_index(idx, p)
    pickle.dump(list(range(1000)), open(p + ".metadata", "wb"))
    lock = threading.Lock()          # rag.py has none; the pickle and the index file are two writes, so the test
    errors, stop = [], []            # serialises (file, pickle) PAIRS -- what is checked is the index object itself

    def uploader():
        try:
            for i in range(10):
                with lock:
                    index = faiss.read_index(p)
                    meta = pickle.load(open(p + ".metadata", "rb"))
                n0 = index.ntotal
                index.add(x[n0:n0 + 300])            # outside the lock: other requests hold views meanwhile
                with lock:
                    faiss.write_index(index, p)
                    pickle.dump(meta + list(range(n0, n0 + 300)), open(p + ".metadata", "wb"))
        except Exception as e:       # noqa: BLE001
            errors.append(repr(e))
        stop.append(1)

    def searcher(seed):
        try:
            q = O.synth(seed, 0, 1, d)
            while not stop:
                with lock:
                    index = faiss.read_index(p)
                    meta = pickle.load(open(p + ".metadata", "rb"))
                n = index.ntotal
                assert n == len(meta), (n, len(meta))
                D, I = index.search(q, 5)
                assert I.max() < n and I.min() >= 0
                [meta[i] for i in I[0]]                # rag.py:118-119
                Dr, Ir = O.search(x[:n], q, 5)
                assert np.array_equal(I, Ir), (n, I, Ir)
        except Exception as e:       # noqa: BLE001
            errors.append(repr(e))

    ts = [threading.Thread(target=uploader)] + [threading.Thread(target=searcher, args=(100 + i,)) for i in range(3)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors[:3]
    assert faiss.read_index(p).ntotal == 4000


def test_long_query_batch_is_chunked(faiss):
    """More than 4096 queries go through the tensor-core kernel in chunks (faiss's query block
    size); results must not depend on the chunking."""
    x = O.synth(1234, 0, 30000, 64)
    q = O.synth(5678, 0, 4096 + 300, 64)
    idx = _build(faiss, x)
    D, I = idx.search(q, 5)
    D2, I2 = idx.search(q[4000:], 5)
    assert np.array_equal(I[4000:], I2) and np.array_equal(D[4000:], D2)
    sel = [0, 4095, 4096, 4395]
    assert_knn_matches(D[sel], I[sel], x, q[sel], 5, expanded=True, what="chunked")


def test_adversarial_row_order(faiss):
    """Rows ordered so that every later row is closer to every query than all earlier ones: the
    running thresholds never help, every tile floods the candidate lists and every list is
    compacted again and again.  Slow by construction, but the answer must still be exact."""
    n, d = 60000, 32
    u = np.zeros(d, np.float32)
    u[0] = 1.0
    scale = (2.0 - np.arange(n, dtype=np.float32) / n)[:, None]        # 2.0 -> 1.0, strictly decreasing
    x = (scale * u[None, :]).astype(np.float32)
    x[:, 1:] = O.synth(9, 0, n, d - 1) * 1e-3
    q = O.synth(10, 0, 40, d) * 1e-2                                    # near the origin
    for path in (2, 1):
        idx = _build(faiss, x)
        idx.set_option("path", path)
        D, I = idx.search(q, 10)
        assert_knn_matches(D, I, x, q, 10, expanded=(path == 2), what="adversarial path=%d" % path)
        assert I.min() >= n - 200                                       # the neighbours are the last rows


def test_concurrent_searches_from_threads(faiss):
    """Flask's dev server is threaded (rag.py:204) and builds an index object per request
    (rag.py:153): independent handles must not interfere, and concurrent searches on ONE handle
    (the resident-index cache hands the same object to every request) must be safe."""
    import threading
    x = O.synth(1234, 0, 50000, 128)
    shared = _build(faiss, x)
    qs = [O.synth(100 + t, 0, 3 if t % 2 else 40, 128) for t in range(8)]
    refs = [O.search(x, q, 5) for q in qs]
    errors = []

    def worker(t):
        try:
            own = _build(faiss, x[: 20000 + 1000 * t])
            for _ in range(5):
                D, I = shared.search(qs[t], 5)
                assert_knn_matches(D, I, x, qs[t], 5, refs[t][0], refs[t][1], expanded=qs[t].shape[0] >= 20, what="shared t=%d" % t)
                D, I = own.search(qs[t], 5)
                assert_knn_matches(D, I, x[: 20000 + 1000 * t], qs[t], 5, expanded=qs[t].shape[0] >= 20, what="own t=%d" % t)
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors


@pytest.mark.parametrize("n", [1, 5, 255, 256, 257, 1000])
@pytest.mark.parametrize("d", [4, 36, 128])
def test_tensor_path_small_and_ragged(faiss, n, d):
    """Batched (tensor-core) search on tiny and ragged shapes: fewer rows than one tile, fewer
    rows than k (padding), row and column counts that are not multiples of the tile."""
    x = O.synth(41, 0, n, d)
    q = O.synth(42, 0, 21, d)
    idx = _build(faiss, x)
    D, I = idx.search(q, 10)
    assert idx.stats()["path"] == "tensor"
    Dr, Ir = O.search(x, q, 10)
    assert np.array_equal(I < 0, Ir < 0)
    assert_knn_matches(D, I, x, q, 10, Dr, Ir, expanded=True, what="n=%d d=%d" % (n, d))


def test_bf16_screening_incremental_adds_and_fallbacks(faiss):
    """bf16 screening keeps bf16 copies of the rows: rows added AFTER the first bf16 search (and
    after a capacity growth) must be searchable, a reset must drop the copies, d % 8 != 0 is served
    by the TF32 screening, and an overflowing slack zone still ends in the 3xTF32 rerun."""
    x = O.synth(61, 0, 30000, 128)
    q = O.synth(62, 0, 40, 128)
    idx = faiss.IndexFlatL2(128)
    idx.set_option("tensor_mode", 3)
    idx.add(x[:5000])
    D, I = idx.search(q, 20)
    assert idx.stats()["tensor_mode"] == "bf16screen+rerank"
    assert_knn_matches(D, I, x[:5000], q, 20, expanded=True, what="bf16 first")
    idx.add(x[5000:])                                     # grows the row storage and the bf16 copies
    x2 = x.copy()
    D, I = idx.search(q, 20)
    assert_knn_matches(D, I, x2, q, 20, expanded=True, what="bf16 after growth")
    big = (x[:64] * 300.0).astype(np.float32)             # larger norms and rounding errors arrive later
    idx.add(big)
    D, I = idx.search(q, 20)
    assert_knn_matches(D, I, np.concatenate([x2, big]), q, 20, expanded=True, what="bf16 after large rows")
    idx.reset()
    idx.add(x[:3000])
    D, I = idx.search(q, 20)
    assert idx.stats()["tensor_mode"] == "bf16screen+rerank"
    assert_knn_matches(D, I, x[:3000], q, 20, expanded=True, what="bf16 after reset")

    y = O.synth(63, 0, 4000, 36)                          # 36 % 8 != 0: no 16-byte bf16 pitch
    idx = _build(faiss, y)
    idx.set_option("tensor_mode", 3)
    qy = O.synth(64, 0, 32, 36)
    D, I = idx.search(qy, 5)
    assert idx.stats()["tensor_mode"] == "screen+rerank"
    assert_knn_matches(D, I, y, qy, 5, expanded=True, what="bf16 -> tf32 screening")

    rng = np.random.default_rng(0)
    base = rng.standard_normal((1, 64)).astype(np.float32)
    z = np.repeat(base, 20000, 0) + rng.standard_normal((20000, 64)).astype(np.float32) * 1e-6
    z[::1000] = rng.standard_normal((20, 64)).astype(np.float32)
    qz = rng.standard_normal((30, 64)).astype(np.float32)
    idx = _build(faiss, z)
    idx.set_option("tensor_mode", 3)
    D, I = idx.search(qz, 10)
    st = idx.stats()
    assert st["screen_fallbacks"] >= 1 and (st["path"] == "scan" or st.get("tensor_mode") == "3xtf32screen+rerank")
    assert_knn_matches(D, I, z, qz, 10, expanded=True, what="bf16 overflow fallback")


def _scan_fallback(idx):
    import ctypes as C
    from b200knn import _lib
    v = C.c_double()
    assert _lib.lib().b2k_get_stat(idx._h, b"scan_fallback", C.byref(v)) == 0
    return int(v.value)


@pytest.mark.parametrize("n,d,nq,k", [(30000, 64, 1, 5), (50000, 384, 1, 100), (20011, 128, 3, 10), (40000, 256, 19, 37),
                                      (9000, 1024, 2, 496), (300, 64, 1, 10), (5, 8, 4, 3)])
def test_scan_over_bf16_copies_is_exact(faiss, n, d, nq, k):
    """nq < 20 on the bf16 copies of the rows (`scan_rows` = 2): candidates by |q - bf16(x)|^2, exact
    fp32 re-rank -- the same (D, I) as the oracle and as the fp32-row scan (both end in a direct fp32
    sum; only the summation order differs), on N(0,1) rows and on unit-norm rows with self matches."""
    rng = np.random.default_rng(n + k)
    for unit in (False, True):
        x = O.synth(71, 0, n, d)
        q = O.synth(72, 0, nq, d)
        if unit:
            x, q = unit_rows(x), unit_rows(q)
            q[0] = x[rng.integers(0, n)]
        idx = _build(faiss, x)
        idx.set_option("path", 1)
        idx.set_option("scan_rows", 2)
        D, I = idx.search(q, k)
        st = idx.stats()
        assert st["path"] == "scan" and st["scan_rows"] == "bf16+rerank"
        assert _scan_fallback(idx) == 0
        if unit:
            assert D[0, 0] == 0.0
        assert_knn_matches(D, I, x, q, k, what="bf16 scan n=%d d=%d nq=%d k=%d unit=%s" % (n, d, nq, k, unit))
        ref = _build(faiss, x)
        ref.set_option("path", 1)
        ref.set_option("scan_rows", 1)
        Dr, Ir = ref.search(q, k)
        assert ref.stats()["scan_rows"] == "fp32"
        assert_knn_matches(D, I, x, q, k, Dr, Ir, what="bf16 scan vs fp32 scan")


def test_scan_over_bf16_copies_falls_back_and_follows_adds(faiss):
    """Thousands of rows at (almost) the k-th distance do not fit the candidate list: the
    conditional fp32 scan behind the re-rank must redo the search (flag raised, results still
    exact).  Rows added after the bf16 copies exist are found; d % 8 != 0 and k > 496 use fp32."""
    rng = np.random.default_rng(3)
    base = rng.standard_normal((1, 64)).astype(np.float32)
    z = np.repeat(base, 20000, 0) + rng.standard_normal((20000, 64)).astype(np.float32) * 1e-6
    z[::1000] = rng.standard_normal((20, 64)).astype(np.float32)
    qz = rng.standard_normal((2, 64)).astype(np.float32)
    idx = _build(faiss, z)
    idx.set_option("scan_rows", 2)
    D, I = idx.search(qz, 30)                              # only 20 rows lie outside the cluster: the 30th is inside
    assert idx.stats()["scan_rows"] == "bf16+rerank" and _scan_fallback(idx) == 1
    assert_knn_matches(D, I, z, qz, 30, what="bf16 scan fallback")
    D, I = idx.search(qz, 3)                               # far from the cluster: no fallback needed
    assert _scan_fallback(idx) == 0
    assert_knn_matches(D, I, z, qz, 3, what="bf16 scan, same data, small k")

    x = O.synth(73, 0, 20000, 128)
    q = O.synth(74, 0, 2, 128)
    idx = faiss.IndexFlatL2(128)
    idx.set_option("scan_rows", 2)
    idx.add(x[:6000])
    D, I = idx.search(q, 8)
    assert_knn_matches(D, I, x[:6000], q, 8, what="bf16 scan first")
    idx.add(x[6000:])
    near = (q[:1] + 1e-3).astype(np.float32)               # a new nearest neighbour arrives last
    idx.add(near)
    D, I = idx.search(q, 8)
    assert I[0, 0] == 20000 and _scan_fallback(idx) == 0
    assert_knn_matches(D, I, np.concatenate([x, near]), q, 8, what="bf16 scan after adds")
    D, I = idx.search(q, 600)                              # 2k + 32 > 1024: fp32 rows
    assert idx.stats()["scan_rows"] == "fp32"
    assert_knn_matches(D, I, np.concatenate([x, near]), q, 600, what="large k on fp32 rows")

    y = O.synth(75, 0, 3000, 36)
    idx = _build(faiss, y)
    idx.set_option("scan_rows", 2)
    qy = O.synth(76, 0, 1, 36)
    D, I = idx.search(qy, 5)
    assert idx.stats()["scan_rows"] == "fp32"
    assert_knn_matches(D, I, y, qy, 5, what="d % 8 != 0 on fp32 rows")


def test_screening_chain_stays_exact_when_norms_dwarf_distances(faiss):
    """Rows 100 + N(0,1): squared norms ~1.3e6 against squared distances ~250.  The slack of the bf16
    screen (proportional to |q||x|) swallows the whole database, so the search must climb down its
    chain -- 3xTF32 as a screen with the exact re-rank, else the scan kernel -- and still return the
    true neighbours with their direct-difference distances (the plain 3xTF32 expansion is off by
    ~1 % here, as faiss's own sgemm path would be).  Checked against the float64 ground truth."""
    rng = np.random.default_rng(11)
    n, d, nq, k = 60000, 128, 32, 20
    x = (rng.standard_normal((n, d)) + 100.0).astype(np.float32)
    q = (rng.standard_normal((nq, d)) + 100.0).astype(np.float32)
    idx = _build(faiss, x)
    D, I = idx.search(q, k)
    st = idx.stats()
    assert st["screen_fallbacks"] >= 1
    Dt, It = O.truth_f64(x, q, k)
    dd = O.dist_f64(x, q, I)
    assert np.all(dd <= Dt[:, k - 1:k] * (1 + 4e-6)), "a returned row is farther than the true k-th neighbour"
    assert np.abs(D.astype(np.float64) - dd).max() <= 4e-6 * dd.max()
    assert all(len(set(r.tolist())) == k for r in I)
    # after two overflowing searches in a row the index starts at the 3xTF32 screen by itself
    idx.search(q, k)
    before = idx.stats()["screen_fallbacks"]
    D3, I3 = idx.search(q, k)
    st3 = idx.stats()
    # (no cheap attempt any more: at most the 3xTF32 screen itself overflows for a crowded query and the scan answers)
    assert st3["screen_fallbacks"] - before <= 1
    assert st3.get("tensor_mode") == "3xtf32screen+rerank" or st3["path"] == "scan"
    assert np.array_equal(I3, I)
    np.testing.assert_allclose(D3, D, rtol=1e-6)           # (the re-rank and the scan sum in different orders)
    idx.set_option("tensor_mode", 4)                       # the second link of the chain on its own
    D4, I4 = idx.search(q, k)
    assert idx.stats().get("tensor_mode") == "3xtf32screen+rerank" or idx.stats()["path"] == "scan"
    dd4 = O.dist_f64(x, q, I4)
    assert np.all(dd4 <= Dt[:, k - 1:k] * (1 + 4e-6))


def test_preallocated_and_pinned_outputs(faiss):
    """faiss's optional D / I output arguments, pageable and page-locked (the latter are copied to
    directly, without the staging buffer): same answers as the plain call, outputs returned."""
    import torch
    x = O.synth(81, 0, 30000, 64)
    idx = _build(faiss, x)
    for nq, k in ((3, 7), (40, 10)):
        q = O.synth(82, 0, nq, 64)
        Dref, Iref = idx.search(q, k)
        D = np.empty((nq, k), np.float32); I = np.empty((nq, k), np.int64)
        D2, I2 = idx.search(q, k, D=D, I=I)
        assert D2 is D and I2 is I and np.array_equal(D, Dref) and np.array_equal(I, Iref)
        qp = torch.from_numpy(q).pin_memory().numpy()
        Dp = torch.empty((nq, k), dtype=torch.float32, pin_memory=True).numpy()
        Ip = torch.empty((nq, k), dtype=torch.int64, pin_memory=True).numpy()
        idx.search(qp, k, D=Dp, I=Ip)
        assert np.array_equal(Dp, Dref) and np.array_equal(Ip, Iref)
        qd = torch.from_numpy(q).cuda()
        Dd = torch.empty((nq, k), dtype=torch.float32, device="cuda"); Id = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        idx.search(qd, k, D=Dd, I=Id)
        assert np.array_equal(Dd.cpu().numpy(), Dref) and np.array_equal(Id.cpu().numpy(), Iref)
    with pytest.raises(AssertionError):
        idx.search(O.synth(82, 0, 3, 64), 7, D=np.empty((3, 8), np.float32), I=np.empty((3, 7), np.int64))


def _hard_cases(rng, n, d, nq):
    g = rng.standard_normal((n, d)).astype(np.float32)
    s = np.exp(rng.standard_normal((n, 1)) * 1.5).astype(np.float32)
    yield "heavy-tailed row norms", g * s, rng.standard_normal((nq, d)).astype(np.float32)
    c = rng.standard_normal((200, d)).astype(np.float32)
    cl = (c[rng.integers(0, 200, n)] + rng.standard_normal((n, d)).astype(np.float32) * 1e-3).astype(np.float32)
    yield "tight clusters", cl, (c[:nq] + rng.standard_normal((nq, d)).astype(np.float32) * 1e-3).astype(np.float32)
    u = g / np.linalg.norm(g, axis=1, keepdims=True)
    qs = u[rng.integers(0, n, nq)] + rng.standard_normal((nq, d)).astype(np.float32) * 0.01
    yield "unit norm, queries next to stored rows", u.astype(np.float32), (qs / np.linalg.norm(qs, axis=1, keepdims=True)).astype(np.float32)
    yield "tiny magnitudes", (g * 1e-18).astype(np.float32), (rng.standard_normal((nq, d)) * 1e-18).astype(np.float32)
    lat = rng.integers(-3, 4, (n, d)).astype(np.float32)
    yield "integer lattice, many exact ties", lat, rng.integers(-3, 4, (nq, d)).astype(np.float32)


@pytest.mark.parametrize("opts", [{"path": 2, "tensor_mode": 3}, {"path": 2, "tensor_mode": 2}, {"path": 1, "scan_rows": 2}],
                         ids=["tensor-bf16-screen", "tensor-tf32-screen", "scan-bf16"])
def test_screens_are_exact_on_hard_distributions(faiss, opts):
    """The reduced-precision screens against the float64 ground truth on distributions chosen to hurt
    them (scripts/stress_exact.py has the long version): every returned row must be at least as close
    as the true k-th neighbour, ids unique, distances = the direct fp32 sums."""
    rng = np.random.default_rng(5)
    n, d, k = 50000, 128, 20
    nq = 3 if opts["path"] == 1 else 48
    for name, x, q in _hard_cases(rng, n, d, nq):
        idx = _build(faiss, x)
        for kk, v in opts.items():
            idx.set_option(kk, v)
        D, I = idx.search(q, k)
        Dt, It = O.truth_f64(x, q, k)
        dd = O.dist_f64(x, q, I)
        kth = Dt[:, k - 1:k]
        assert np.all(dd <= kth * (1 + 4e-6) + 1e-37), name
        rel = np.where(dd > 1e-37, np.abs(D.astype(np.float64) - dd) / np.maximum(dd, 1e-300), 0.0)
        assert rel.max() < 4e-6, (name, rel.max())
        assert all(len(set(r.tolist())) == k for r in I), name


def test_nonfinite_rows_and_queries_on_the_tensor_path(faiss):
    """inf / NaN rows are never selected and do not disturb their neighbours in the tile; a NaN
    query gets all padding (no distance is < FLT_MAX), the other queries are unaffected."""
    x = O.synth(51, 0, 3000, 64)
    x[5, 3] = np.inf
    x[17, 0] = np.nan
    x[300] = np.inf
    q = O.synth(52, 0, 24, 64)
    q[7, 1] = np.nan
    for mode in (2, 1, 3, 4):
        idx = _build(faiss, x)
        idx.set_option("tensor_mode", mode)
        D, I = idx.search(q, 8)
        assert idx.stats()["path"] == "tensor"
        assert not np.isin(I, [5, 17, 300]).any()
        assert np.all(I[7] == -1) and np.all(D[7] == O.FLT_MAX)
        ok = [i for i in range(24) if i != 7]
        xs = x.copy()
        xs[[5, 17, 300]] = 1e6                       # far away, finite: same neighbours for the oracle
        assert_knn_matches(D[ok], I[ok], xs, q[ok], 8, expanded=True, what="nonfinite mode=%d" % mode)


def test_device_adds_are_ordered_before_growth_write_and_readback(faiss, tmp_path):
    """Rows added from a CUDA tensor land on the caller's stream; a later add that regrows the
    storage, write_index and reconstruct_n run on the handle's own stream and must wait for them
    (rag.py:80 add -> :82 write_index, with the rows coming from a device-side encoder)."""
    import torch
    d = 256
    side = torch.cuda.Stream()
    idx = faiss.IndexFlatL2(d)
    a = faiss.synth(31, 0, 400_000, d)
    b = faiss.synth(31, 400_000, 700_000, d)
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        # keep the side stream busy so the row copy is still pending when the host moves on
        junk = torch.empty(1 << 28, dtype=torch.float32, device="cuda")
        for _ in range(4):
            junk.normal_()
        idx.add(a)                       # on-device add, enqueued on `side`
    idx.add(b.cpu().numpy())             # host add on the handle's stream: grows (capacity 400k -> 1.1M)
    path = str(tmp_path / "ordered.index")
    faiss.write_index(idx, path)
    got = idx.reconstruct_n(0, idx.ntotal)
    want = O.synth(31, 0, 1_100_000, d)
    assert np.array_equal(got, want)
    dd, rows = O.read_ixf2(path)
    assert dd == d and np.array_equal(rows, want)
    del junk


def test_misaligned_device_queries(faiss):
    """A torch view whose storage offset is not a multiple of 16 bytes is still a valid query batch."""
    import torch
    d, n, k = 64, 600_000, 10      # 600k x 64 = 38M elements: the bf16 scan + re-rank serves nq < 20
    idx = faiss.IndexFlatL2(d)
    x = O.synth(3, 0, n, d)
    idx.add(x)
    for nq in (2, 40):
        q = O.synth(4, 0, nq, d)
        buf = torch.empty(nq * d + 1, dtype=torch.float32, device="cuda")
        qv = buf[1:].view(nq, d)
        qv.copy_(torch.from_numpy(q))
        assert qv.data_ptr() % 16 != 0
        D, I = idx.search(qv, k)
        assert_knn_matches(D.cpu().numpy(), I.cpu().numpy(), x, q, k, expanded=nq >= 20, what="misaligned nq=%d" % nq)


def test_zero_and_far_queries_on_the_screening_path(faiss):
    """The screening slack must also cover the fp32 rounding of the approximate distance itself: with
    q = 0, or |q| far beyond every row, the product term of the slack vanishes or is dwarfed."""
    rng = np.random.default_rng(5)
    n, d, k = 200_000, 64, 10
    x = rng.standard_normal((n, d)).astype(np.float32)
    x[::7] *= 1.0 + 1e-7 * rng.standard_normal((len(x[::7]), 1)).astype(np.float32)   # near-ties in norm
    q = rng.standard_normal((32, d)).astype(np.float32)
    q[0] = 0.0
    q[1] = 1e3
    q[2] = -3e4
    idx = _build(faiss, x)
    for mode in (3, 2, 4):
        idx.set_option("path", 2)
        idx.set_option("tensor_mode", mode)
        D, I = idx.search(q, k)
        Dt, It = O.truth_f64(x, q, k)
        d64 = O.dist_f64(x, q, I)
        assert np.all(np.abs(d64 - Dt) <= 1e-5 * np.maximum(Dt, 1e-30) + 1e-30), "mode %d: a nearer row was missed" % mode
        assert_knn_matches(D, I, x, q, k, expanded=True, what="zero/far queries, mode %d" % mode)


def test_prepare_builds_the_copies_before_the_first_search(faiss):
    import torch
    d, n = 128, 1_000_000
    idx = faiss.IndexFlatL2(d)
    x = O.synth(8, 0, n, d)
    idx.add(x)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    idx.prepare()
    free1 = torch.cuda.mem_get_info(0)[0]
    assert free0 - free1 >= n * d * 2 * 0.9            # the bf16 copies (+ 32 bytes of norm record per row) exist now
    q = O.synth(9, 0, 64, d)
    D, I = idx.search(q, 10)
    free2 = torch.cuda.mem_get_info(0)[0]
    assert free1 - free2 < n * d                        # the search allocated its workspace only (~100 MB of candidate lists)
    assert_knn_matches(D, I, x, q, 10, expanded=True, what="after prepare")


def test_a_few_huge_rows_do_not_widen_every_querys_slack(faiss):
    """A handful of rows with 30 times the typical norm (and a later add that brings more): the batched search keeps
    its bf16 screen -- the rows are taken out of the screen and re-ranked exactly for every query -- and stays exact,
    also for queries that sit on or next to such a row."""
    rng = np.random.default_rng(11)
    n, d, nq, k = 700_000, 64, 512, 10
    x = rng.standard_normal((n, d)).astype(np.float32)
    big = rng.choice(n, 20, replace=False)
    x[big] *= 30.0
    q = rng.standard_normal((nq, d)).astype(np.float32)
    q[0] = x[big[0]]                                   # a query ON an outlier row
    q[1] = x[big[1]] + 0.5 * rng.standard_normal(d).astype(np.float32)
    q[2] = 25.0 * rng.standard_normal(d).astype(np.float32)   # a far-out query: the outliers are its neighbours
    idx = _build(faiss, x)
    D, I = idx.search(q, k)
    st = idx.stats()
    assert st["path"] == "tensor" and st["tensor_mode"] == "bf16screen+rerank" and st["screen_fallbacks"] == 0, st
    assert I[0, 0] == big[0] and D[0, 0] == 0.0
    Dt, It = O.truth_f64(x, q[:48], k)          # (float64 brute force: a sample of the batch, the special queries included)
    d64 = O.dist_f64(x, q[:48], I[:48])
    # (the float64 truth uses the norm expansion: a self-match comes out as ~1e-11, hence the absolute term)
    assert np.all(np.abs(d64 - Dt) <= 1e-5 * Dt + 1e-6), "a nearer row was missed"
    assert_knn_matches(D, I, x, q, k, expanded=True, what="outlier rows")
    # more rows, among them new outliers (and the old ones stay): the set is recomputed after the add
    y = rng.standard_normal((5000, d)).astype(np.float32)
    y[::500] *= 40.0
    idx.add(y)
    xa = np.concatenate([x, y])
    q[3] = y[500]
    D, I = idx.search(q, k)
    st = idx.stats()
    assert st["tensor_mode"] == "bf16screen+rerank" and st["screen_fallbacks"] == 0, st
    assert I[3, 0] == n + 500 and D[3, 0] == 0.0
    Dt, It = O.truth_f64(xa, q[:48], k)
    d64 = O.dist_f64(xa, q[:48], I[:48])
    assert np.all(np.abs(d64 - Dt) <= 1e-5 * Dt + 1e-6), "a nearer row was missed after the add"
    assert_knn_matches(D, I, xa, q, k, expanded=True, what="outlier rows after the add")
    # too many large rows to take out (more than 64): nothing is taken out, the search is still exact (by whatever path)
    z = rng.standard_normal((300_000, d)).astype(np.float32)
    z[::1000] *= 30.0
    idx2 = _build(faiss, z)
    qz = rng.standard_normal((300, d)).astype(np.float32)
    D, I = idx2.search(qz, k)
    assert_knn_matches(D, I, z, qz, k, expanded=True, what="many large rows")
    # a view that ends before some outlier rows must not return them
    v = idx._view()
    idx.add(40.0 * rng.standard_normal((3, d)).astype(np.float32))
    D, I = idx.search(q[:200], k)
    Dv, Iv = v.search(q[:200], k)
    assert Iv.max() < n + 5000 and I.max() < n + 5003
    assert_knn_matches(Dv, Iv, xa, q[:200], k, expanded=True, what="view below later outliers")
