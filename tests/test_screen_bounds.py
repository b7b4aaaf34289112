"""The inequalities the reduced-precision screens of libb200knn rest on, checked on the CPU with
numpy float64 (no GPU, no library): if these hold, a screen that keeps every row inside its slack
zone cannot lose a true neighbour.  DESIGN.md section 4 states them; csrc/k2_scan.cu (bf16 scan) and
csrc/k3_tc.cu (bf16 / TF32 tensor screens) implement them."""
import numpy as np
import pytest
import torch

from oracle import oracle as O


def _bf16(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def _tf32_trunc(a):
    return (np.ascontiguousarray(a, dtype=np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


def _datasets(rng, n, d, nq):
    g = rng.standard_normal((n, d)).astype(np.float32)
    yield "gaussian", g, rng.standard_normal((nq, d)).astype(np.float32)
    yield "heavy-tailed norms", (g * np.exp(rng.standard_normal((n, 1)) * 1.5)).astype(np.float32), rng.standard_normal((nq, d)).astype(np.float32)
    yield "offset mean", (g + 100.0).astype(np.float32), (rng.standard_normal((nq, d)) + 100.0).astype(np.float32)
    u = g / np.linalg.norm(g, axis=1, keepdims=True)
    yield "unit norm, near-duplicates", u.astype(np.float32), (u[:nq] + 1e-3 * rng.standard_normal((nq, d))).astype(np.float32)
    yield "tiny magnitudes", (g * 1e-18).astype(np.float32), (rng.standard_normal((nq, d)) * 1e-18).astype(np.float32)


@pytest.mark.parametrize("k", [1, 10, 100])
def test_scan_zone_contains_every_true_neighbour(k):
    """bf16 scan (K2): d~ = |q - bf16(x)|^2, E = max |x - bf16(x)|.  Every exact top-k row has
    sqrt(d~) <= sqrt(T~) + 2E, T~ = k-th smallest d~ (triangle inequality)."""
    rng = np.random.default_rng(k)
    for name, x, q in _datasets(rng, 4000, 64, 6):
        xb = _bf16(x).astype(np.float64)
        x64, q64 = x.astype(np.float64), q.astype(np.float64)
        E = np.sqrt(((x64 - xb) ** 2).sum(1)).max()
        for qi in q64:
            d_exact = ((qi[None, :] - x64) ** 2).sum(1)
            d_apx = ((qi[None, :] - xb) ** 2).sum(1)
            T = np.sort(d_apx)[k - 1]
            top = np.argsort(d_exact, kind="stable")[:k]
            assert np.all(np.sqrt(d_apx[top]) <= (np.sqrt(T) + 2 * E) * (1 + 1e-12)), name


@pytest.mark.parametrize("rounding", ["bf16", "tf32"])
def test_tensor_screen_slack_bounds_the_product_error(rounding):
    """Tensor screens (K3): |q.x - q~.x~| <= |q - q~| max|x| + |q~| max|x - x~| (measured rounding
    errors, the bf16 slack) and <= 2^-9 |q||x| for operands truncated to TF32; hence every exact
    top-k row has d~ <= T~ + 2 sigma with sigma twice that bound."""
    rng = np.random.default_rng(3)
    k = 10
    for name, x, q in _datasets(rng, 3000, 96, 5):
        rnd = _bf16 if rounding == "bf16" else _tf32_trunc
        xr, qr = rnd(x).astype(np.float64), rnd(q).astype(np.float64)
        x64, q64 = x.astype(np.float64), q.astype(np.float64)
        nx = np.sqrt((x64 ** 2).sum(1))
        ex = np.sqrt(((x64 - xr) ** 2).sum(1))
        for qi, qri in zip(q64, qr):
            err = np.abs(x64 @ qi - xr @ qri)
            if rounding == "bf16":
                bound = np.linalg.norm(qi - qri) * nx.max() + np.linalg.norm(qri) * ex.max()
            else:
                bound = 2.0 ** -9 * np.linalg.norm(qi) * nx.max()
            assert np.all(err <= bound * (1 + 1e-9) + 1e-300), name
            sigma = 2 * bound
            d_exact = ((qi[None, :] - x64) ** 2).sum(1)
            d_apx = (qi ** 2).sum() + (x64 ** 2).sum(1) - 2 * (xr @ qri)
            T = np.sort(d_apx)[k - 1]
            top = np.argsort(d_exact, kind="stable")[:k]
            assert np.all(d_apx[top] <= T + 2 * sigma + 1e-9 * abs(T)), name


def test_oracle_agrees_with_screen_then_rerank_emulation():
    """Emulated bf16 scan + exact re-rank (2k + 32 candidates, as the library picks them) returns the
    oracle's neighbours whenever the coverage check passes."""
    rng = np.random.default_rng(9)
    x = rng.standard_normal((5000, 32)).astype(np.float32)
    q = rng.standard_normal((4, 32)).astype(np.float32)
    k, kp = 10, 52
    xb = _bf16(x).astype(np.float64)
    E = np.sqrt(((x.astype(np.float64) - xb) ** 2).sum(1)).max()
    Dref, Iref = O.search(x, q, k)
    for i, qi in enumerate(q.astype(np.float64)):
        d_apx = ((qi[None, :] - xb) ** 2).sum(1)
        cand = np.argsort(d_apx, kind="stable")[:kp]
        zone = (np.sqrt(d_apx[cand[k - 1]]) + 2 * E) ** 2
        assert d_apx[cand[kp - 1]] > zone          # coverage check passes on this data
        d_c = ((qi[None, :] - x[cand].astype(np.float64)) ** 2).sum(1)
        best = cand[np.lexsort((cand, d_c))[:k]]
        assert np.array_equal(best, Iref[i])
