"""CPU tests of the oracle itself: against the float64 ground truth, hand-derived known answers
and the committed golden fixtures.  The known answers marked [faiss-upstream] come from the
recorded behaviour of faiss and are UNVERIFIED against a real faiss build:
faiss-cpu is not installable here, so parity with the reference is unpinned."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from util import unit_rows

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("d", [1, 3, 17, 384, 1027])
@pytest.mark.parametrize("nq", [1, 19, 20, 33])
def test_oracle_vs_float64(d, nq):
    rng = np.random.default_rng(d * 100 + nq)
    x = rng.standard_normal((1500, d)).astype(np.float32)
    q = rng.standard_normal((nq, d)).astype(np.float32)
    k = 10
    D, I = O.search(x, q, k)
    Dt, It = O.truth_f64(x, q, k)
    scale = np.maximum(Dt, (q.astype(np.float64) ** 2).sum(1)[:, None] + (x.astype(np.float64) ** 2).sum(1)[It])
    assert np.all(np.abs(D - Dt) <= 1e-5 * scale)
    # ids may differ only on float32 near-ties
    mism = I != It
    if mism.any():
        d_a = O.dist_f64(x, q, I)
        assert np.all(np.abs(d_a - Dt)[mism] <= 2e-5 * scale[mism])


def test_direct_and_expansion_agree_on_lattice():
    """Small-integer vectors: every distance is exactly representable, so both numerics, every
    summation order and the GPU must give bit-identical D and I."""
    rng = np.random.default_rng(7)
    x = rng.integers(-8, 9, size=(3000, 48)).astype(np.float32)
    q = rng.integers(-8, 9, size=(25, 48)).astype(np.float32)
    D1, I1 = O.search(x, q, 20, O.MODE_DIRECT)
    D2, I2 = O.search(x, q, 20, O.MODE_EXPAND)
    D3, I3 = O.search(x, q, 20, O.MODE_DIRECT_ROWPAR)
    Dt, It = O.truth_f64(x, q, 20)
    assert np.array_equal(D1, D2) and np.array_equal(I1, I2)
    assert np.array_equal(D1, D3) and np.array_equal(I1, I3)
    assert np.array_equal(D1.astype(np.float64), Dt) and np.array_equal(I1, It)


def test_padding_when_fewer_rows_than_k():
    """[faiss-upstream] ntotal < k pads with I=-1, D=FLT_MAX; rag.py:118-119 then indexes
    metadata[-1] (0 < ntotal < 5) or raises IndexError (ntotal == 0)."""
    x = np.array([[0, 0], [3, 4], [1, 0]], np.float32)
    q = np.array([[0, 0]], np.float32)
    D, I = O.search(x, q, 5)
    assert I.tolist() == [[0, 2, 1, -1, -1]]
    assert D[0, :3].tolist() == [0.0, 1.0, 25.0]
    assert np.all(D[0, 3:] == np.float32(3.4028235e38))
    D, I = O.search(np.zeros((0, 2), np.float32), q, 5)
    assert I.tolist() == [[-1] * 5] and np.all(D == np.float32(3.4028235e38))


def test_ties_resolve_to_lower_ids():
    """[faiss-upstream] strict '<' against the current worst + scan in id order: among equal
    distances the lower ids stay, output ordered by (distance, id)."""
    x = np.ones((50, 8), np.float32)
    q = np.zeros((1, 8), np.float32)
    for mode in (O.MODE_DIRECT, O.MODE_EXPAND):
        D, I = O.search(x, np.repeat(q, 21, 0), 7, mode)
        assert I[0].tolist() == list(range(7)) and np.all(D == 8.0)


def test_query_equal_to_row_gives_zero():
    rng = np.random.default_rng(3)
    x = unit_rows(rng.standard_normal((200, 1024)))
    q = x[[5, 17]].copy()
    D, I = O.search(x, q, 3, O.MODE_DIRECT)
    assert I[:, 0].tolist() == [5, 17] and np.all(D[:, 0] == 0.0)
    D, I = O.search(x, np.repeat(q, 10, 0), 3, O.MODE_EXPAND)
    assert np.all(D >= 0.0) and np.all(D[:, 0] <= 1e-6)


def test_nonfinite_rows_are_never_selected():
    x = np.array([[1.0], [np.inf], [np.nan], [2.0]], np.float32)
    D, I = O.search(x, np.zeros((1, 1), np.float32), 4)
    assert I.tolist() == [[0, 3, -1, -1]]


def test_merge_matches_single_index():
    rng = np.random.default_rng(11)
    x = rng.standard_normal((4000, 32)).astype(np.float32)
    q = rng.standard_normal((6, 32)).astype(np.float32)
    k = 10
    D, I = O.search(x, q, k)
    parts = [O.search(x[a:b], q, k) for a, b in ((0, 1000), (1000, 1003), (1003, 4000))]
    Dl = np.stack([p[0] for p in parts])
    Il = np.stack([np.where(p[1] >= 0, p[1] + off, -1) for p, off in zip(parts, (0, 1000, 1003))])
    Dm, Im = O.merge(Dl, Il)
    assert np.array_equal(Dm, D) and np.array_equal(Im, I)


def test_synth_stream_is_stable_and_unit_variance():
    a = O.synth(1234, 0, 4096, 96)
    b = O.synth(1234, 1000, 10, 96)
    assert np.array_equal(a[1000:1010], b)
    assert abs(a.mean()) < 0.01 and abs(a.std() - 1.0) < 0.01
    # known answers pin the generator (shared bit-for-bit with the CUDA generator)
    gold = np.load(os.path.join(GOLDEN, "synth_kat.npz"))
    assert np.array_equal(O.synth(int(gold["seed"]), int(gold["row0"]), gold["rows"].shape[0], gold["rows"].shape[1]), gold["rows"])


@pytest.mark.parametrize("name", ["c1_2048x384_nq1_k5", "app_64x1024_unit_nq1_k5", "batch_3000x96_nq32_k100"])
def test_golden_fixtures(name):
    """Fixtures made by tests/golden/make_golden.py: inputs from the synthetic stream, expected
    (D, I) from the float64 brute force.  They guard the oracle against regressions; they are
    NOT outputs of the reference (none can be produced here)."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    x = O.synth(int(g["seed_x"]), 0, int(g["n"]), int(g["d"]))
    q = O.synth(int(g["seed_q"]), 0, int(g["nq"]), int(g["d"]))
    if bool(g["unit"]):
        x, q = unit_rows(x), unit_rows(q)
    D, I = O.search(x, q, int(g["k"]))
    assert np.array_equal(I, g["I"])
    assert np.allclose(D, g["D"], rtol=2e-6, atol=2e-6)


def test_blas_backed_variant_matches_plain_restatement():
    """search_blas (sgemm blocks through torch's MKL, used only as the timed CPU baseline) returns
    the same neighbours as the plain-C expansion path; distances agree to BLAS summation order."""
    x = O.synth(1234, 0, 20000, 96)
    q = O.synth(5678, 0, 33, 96)
    D1, I1 = O.search(x, q, 10, O.MODE_EXPAND)
    D2, I2 = O.search_blas(x, q, 10, db_bs=4096)
    assert np.allclose(D1, D2, rtol=1e-5, atol=1e-4)
    assert (I1 == I2).mean() > 0.99


def test_streaming_search_equals_one_call_and_compare_knn_sees_a_missed_neighbour():
    """O.stream_search (the full-size checker of tests/test_gpu_fullsize.py and bench.py's parity_check)
    gives the answer of one call on all rows under both numerics; O.compare_knn accepts it and
    rejects a result with a wrong neighbour, a wrong distance or wrong padding."""
    d, n, k = 48, 30011, 10
    x = O.synth(7, 0, n, d)
    rows = lambda j0, nb: O.synth(7, j0, nb, d)
    for nq in (3, 25):
        q = O.synth(8, 0, nq, d)
        D, I = O.search(x, q, k)
        Ds, Is = O.stream_search(rows, n, d, q, k, chunk=4096)
        assert np.array_equal(D, Ds) and np.array_equal(I, Is)
        ok = O.compare_knn(D, I, Ds, Is, O.synth_rows(7, d), q, expanded=nq >= 20)
        assert ok["ok"] and ok["rows_with_other_ids"] == 0
        I2 = I.copy()
        I2[1, 4] = next(r for r in range(n) if r not in set(I[1].tolist()))
        assert not O.compare_knn(D, I2, Ds, Is, O.synth_rows(7, d), q, expanded=nq >= 20)["ok"]
        D2 = D.copy()
        D2[0, 0] *= 1.001
        assert not O.compare_knn(D2, I, Ds, Is, O.synth_rows(7, d), q, expanded=nq >= 20)["ok"]
    Dp, Ip = O.stream_search(rows, 4, d, O.synth(8, 0, 2, d), k, chunk=3)
    assert (Ip[:, 4:] == -1).all() and (Dp[:, 4:] == O.FLT_MAX).all()
    assert O.compare_knn(Dp, Ip, Dp, Ip, O.synth_rows(7, d), O.synth(8, 0, 2, d))["ok"]
    Ibad = Ip.copy(); Ibad[0, 3] = -1
    assert not O.compare_knn(Dp, Ibad, Dp, Ip, O.synth_rows(7, d), O.synth(8, 0, 2, d))["ok"]


def test_oracle_agrees_with_sklearn_brute_force():
    """A third implementation the builder did not write: scikit-learn's brute-force
    NearestNeighbors with the squared-euclidean metric, on unit-norm rows (the rag.py:55 distribution)
    and on N(0,1) rows.  Ids must agree except at float ties; distances within 1e-5 of the magnitudes."""
    sk = pytest.importorskip("sklearn.neighbors")
    rng = np.random.default_rng(3)
    for unit in (False, True):
        x = rng.standard_normal((5000, 96)).astype(np.float32)
        q = rng.standard_normal((40, 96)).astype(np.float32)
        if unit:
            x /= np.linalg.norm(x, axis=1, keepdims=True)
            q /= np.linalg.norm(q, axis=1, keepdims=True)
        nn = sk.NearestNeighbors(n_neighbors=10, algorithm="brute", metric="sqeuclidean").fit(x.astype(np.float64))
        Dk, Ik = nn.kneighbors(q.astype(np.float64))
        for qq, mode in ((q, O.MODE_EXPAND), (q[:7], O.MODE_DIRECT)):
            D, I = O.search(x, qq, 10, mode)
            m = len(qq)
            scale = np.maximum(Dk[:m], (qq.astype(np.float64) ** 2).sum(1)[:, None] + 1.0 if mode == O.MODE_EXPAND else Dk[:m])
            assert np.all(np.abs(D - Dk[:m]) <= 1e-5 * scale)
            same = (I == Ik[:m])
            # where the order differs, the two distances involved are equal at rounding level
            assert np.all(np.abs(D[~same] - Dk[:m][~same]) <= 1e-5 * scale[~same])
            assert same.mean() > 0.99
