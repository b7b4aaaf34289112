"""Oracle-checked parity at the FULL sizes of BASELINE.json's configs.

The row sets do not fit comfortably in host memory next to everything else (C3 15 GB, C5 41 GB),
so the oracle streams them: `O.stream_search` regenerates the synthetic rows one million at a
time on the CPU (an implementation independent of the CUDA generator), searches each chunk with
the rule the whole batch selects (rag.py:116 -> faiss: direct differences below 20 queries, norm
expansion from 20 on) and merges.  The GPU searches the whole batch; a sample of its queries,
spread over every 128-query block of the tensor-core kernel, is compared with the oracle by the
parity rule of tests/util.py (distances within 1e-5 relative; ids equal, or tied within that
tolerance in float64).  A missed neighbour shows up as an id the oracle has and the GPU has not.
"""
import numpy as np
import pytest

from oracle import oracle as O
from util import RTOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def faiss():
    import b200knn
    return b200knn


def _check(D, I, sel, Dref, Iref, seed, d, q, expanded, what):
    r = O.compare_knn(D[sel].cpu().numpy(), I[sel].cpu().numpy(), Dref, Iref, O.synth_rows(seed, d), q[sel], rtol=RTOL,
                      expanded=expanded)
    assert r["ok"], "%s: %s (max rel err %s)" % (what, r["why"], r["max_rel_err"])
    return r


def test_c3_full_size_against_streaming_oracle(faiss):
    """configs[2]: 10M x 384, nq=1024, k=100 -- default mode (bf16 screen + exact fp32 re-rank) and
    the 3xTF32 contraction the north_star names (tensor_mode=1), 32 queries over all 8 query blocks;
    plus nq=1 (scan kernel, direct differences) on the same index."""
    N, d, nq, k = 10_000_000, 384, 1024, 100
    idx = faiss.IndexFlatL2(d)
    idx.reserve(N)
    idx.add_synthetic(N, 1234, 0)
    q_dev = faiss.synth(5678, 0, nq, d)
    q = q_dev.cpu().numpy()
    sel = np.array([b * 128 + j for b in range(8) for j in (0, 41, 90, 127)])
    Dref, Iref = O.stream_search(lambda j0, nb: O.synth(1234, j0, nb, d), N, d, q[sel], k, mode=O.MODE_EXPAND)

    D, I = idx.search(q_dev, k)
    st = idx.stats()
    assert st["path"] == "tensor" and st["tensor_mode"] == "bf16screen+rerank" and st["screen_fallbacks"] == 0
    _check(D, I, sel, Dref, Iref, 1234, d, q, True, "C3 default")

    idx.set_option("tensor_mode", 1)
    D1, I1 = idx.search(q_dev, k)
    assert idx.stats()["tensor_mode"] == "3xtf32"
    _check(D1, I1, sel, Dref, Iref, 1234, d, q, True, "C3 3xTF32")
    idx.set_option("tensor_mode", 0)

    # nq = 1: what rag.py:116 sends; direct differences on both sides -> plain relative tolerance
    Dr1, Ir1 = O.stream_search(lambda j0, nb: O.synth(1234, j0, nb, d), N, d, q[:1], k)
    for rows in (0, 1):   # bf16 copies + exact re-rank (default at this size), then the fp32 rows
        idx.set_option("scan_rows", rows)
        Dq, Iq = idx.search(q_dev[:1].contiguous(), k)
        assert idx.stats()["path"] == "scan"
        _check(Dq, Iq, np.array([0]), Dr1, Ir1, 1234, d, q, False, "C3 nq=1 scan_rows=%d" % rows)


def test_c4_shape_d768_nq256_on_the_tensor_path(faiss):
    """configs[3]'s per-GPU shape (d=768, nq=256, k=100) on 2M rows of one shard: the tensor-core path at d=768."""
    N, d, nq, k = 2_000_000, 768, 256, 100
    idx = faiss.IndexFlatL2(d)
    idx.add_synthetic(N, 1234, 0)
    q_dev = faiss.synth(5678, 0, nq, d)
    q = q_dev.cpu().numpy()
    sel = np.array([b * 128 + j for b in range(2) for j in range(0, 128, 8)])
    Dref, Iref = O.stream_search(lambda j0, nb: O.synth(1234, j0, nb, d), N, d, q[sel], k, mode=O.MODE_EXPAND)
    for mode, name in ((0, "bf16screen+rerank"), (2, "screen+rerank"), (1, "3xtf32")):
        idx.set_option("tensor_mode", mode)
        D, I = idx.search(q_dev, k)
        st = idx.stats()
        assert st["path"] == "tensor" and st["tensor_mode"] == name, st
        _check(D, I, sel, Dref, Iref, 1234, d, q, True, "C4 shape, mode %d" % mode)


def test_c5_full_size_k1024(faiss):
    """configs[4]: 10M x 1024, nq=4096, k=1024 at its real size; 8 queries against the streaming oracle."""
    N, d, nq, k = 10_000_000, 1024, 4096, 1024
    idx = faiss.IndexFlatL2(d)
    idx.reserve(N)
    idx.add_synthetic(N, 1234, 0)
    q_dev = faiss.synth(5678, 0, nq, d)
    q = q_dev.cpu().numpy()
    D, I = idx.search(q_dev, k)
    st = idx.stats()
    assert st["path"] == "tensor" and st["screen_fallbacks"] == 0, st
    sel = np.array([0, 127, 700, 1500, 2048, 3000, 3999, 4095])
    Dref, Iref = O.stream_search(lambda j0, nb: O.synth(1234, j0, nb, d), N, d, q[sel], k, chunk=500_000,
                                 mode=O.MODE_EXPAND)
    _check(D, I, sel, Dref, Iref, 1234, d, q, True, "C5")
