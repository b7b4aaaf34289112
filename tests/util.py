"""Shared helpers for the parity tests (the oracle is the checker, never the thing tested)."""
import numpy as np

from oracle import oracle as O

RTOL = 1e-5  # north_star: "distances within 1e-5 relative"


def unit_rows(a):
    a = np.asarray(a, dtype=np.float32)
    n = np.linalg.norm(a.astype(np.float64), axis=1, keepdims=True)
    n[n == 0] = 1.0
    return (a / n).astype(np.float32)


def assert_knn_matches(D, I, x, q, k, Dref=None, Iref=None, rtol=RTOL, expanded=False, what=""):
    """Check (D, I) against the oracle's (Dref, Iref) and the float64 ground truth.

    - distances: |D - Dref| <= rtol * scale, scale = |Dref| on the direct path and
      max(|Dref|, ||q||^2 + ||x||^2) on the expansion path (the magnitude of the terms being
      cancelled -- SURVEY.md section 7, hard part 3);
    - ids: identical to the oracle's, except where the float64 distances of the ids that differ
      are within that tolerance of each other (ties at rounding level);
    - padding: I == -1 exactly where the oracle pads, with D == FLT_MAX.
    """
    D = np.asarray(D)
    I = np.asarray(I)
    x = np.asarray(x, dtype=np.float32)
    q = np.asarray(q, dtype=np.float32)
    if Dref is None:
        Dref, Iref = O.search(x, q, k)
    assert D.shape == Dref.shape == (q.shape[0], k), (what, D.shape, Dref.shape)
    assert I.dtype == np.int64 and D.dtype == np.float32
    pad = Iref < 0
    assert np.array_equal(I < 0, pad), "%s: padding positions differ" % what
    assert np.all(I[pad] == -1) and np.all(D[pad] == O.FLT_MAX), "%s: padding values" % what
    qn = (q.astype(np.float64) ** 2).sum(1)
    xn = (x.astype(np.float64) ** 2).sum(1) if x.shape[0] else np.zeros(0)
    ok = ~pad
    scale = np.abs(Dref.astype(np.float64))
    if expanded and x.shape[0]:
        mag = qn[:, None] + xn[np.where(pad, 0, Iref)]
        scale = np.maximum(scale, mag)
    err = np.abs(D.astype(np.float64) - Dref.astype(np.float64))
    tol = rtol * scale + 1e-30
    bad = ok & (err > tol)
    assert not bad.any(), "%s: distance mismatch, worst rel %.3g at %s" % (
        what, (err[ok] / np.maximum(scale[ok], 1e-30)).max(), np.argwhere(bad)[:3])
    # sorted ascending
    assert np.all(np.diff(np.where(pad, 1e300, D.astype(np.float64)), axis=1) >= 0), "%s: D not ascending" % what
    # ids
    diff_rows = np.where((I != Iref).any(1))[0]
    for i in diff_rows:
        a, b = I[i][~pad[i]], Iref[i][~pad[i]]
        assert len(set(a.tolist())) == len(a), "%s: duplicate ids in row %d" % (what, i)
        d_a = O.dist_f64(x, q[i:i + 1], a[None, :])[0]
        d_b = O.dist_f64(x, q[i:i + 1], b[None, :])[0]
        t = rtol * np.maximum(np.maximum(d_a, d_b), (qn[i] + xn[a]) if expanded else 0) + 1e-30
        # position by position the true distances must agree within tolerance ...
        assert np.all(np.abs(d_a - d_b) <= 2 * t), "%s: row %d ids differ beyond tie tolerance" % (what, i)
