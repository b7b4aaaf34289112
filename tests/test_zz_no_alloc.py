"""b2k_search_rows_noalloc -- the entry point for callers that must not allocate device memory
(an inference engine has already reserved HBM, or the calling thread cannot afford the
device-wide synchronisation of cudaMalloc).  Its scratch is a fixed array inside the library, so
the checks here are: same answers as the regular entry point (b2k_search on the scan path, i.e.
the same kernel, so the comparison is exact) on two calls in a row that reuse that array -- and,
since that compares the kernel with itself, the same answers as the CPU oracle under the parity
rule of the other tests --, the documented size limit is refused loudly, and the call really
allocates nothing.

Collected last on purpose (file name): a run that stops at the first failure has then already
run every other test.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O
from util import assert_knn_matches

pytestmark = pytest.mark.gpu


def _noalloc_search(L, dev, x, q, k, stream, out=None):
    import torch
    nq = q.shape[0]
    if out is None:
        out = (torch.empty((nq, k), dtype=torch.float32, device=x.device),
               torch.empty((nq, k), dtype=torch.int64, device=x.device))
    D, I = out
    rc = L.b2k_search_rows_noalloc(dev, C.c_void_p(x.data_ptr()), x.shape[0], x.shape[1], C.c_void_p(q.data_ptr()),
                                   nq, k, C.c_void_p(D.data_ptr()), C.c_void_p(I.data_ptr()), stream)
    return rc, D, I


def test_noalloc_entry_point_matches_regular_search():
    import torch
    import b200knn
    from b200knn import _lib
    L = _lib.lib()
    dev = 0
    torch.cuda.set_device(dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    x = b200knn.synth(1234, 0, 300_000, 384, device=dev)
    ref = b200knn.IndexFlatL2(384, device=dev)
    ref.add(x)
    ref.set_option("path", 1)            # the scan kernel: what the no-alloc entry point runs ...
    ref.set_option("scan_rows", 1)       # ... on the fp32 rows (the handle would otherwise scan its bf16 copies and re-rank)

    # first call: also resolves the static array's address on this device (before any capture)
    q1 = b200knn.synth(5678, 0, 1, 384, device=dev)
    rc, D1, I1 = _noalloc_search(L, dev, x, q1, 10, stream)
    assert rc == 0, _lib.last_error()
    Dr, Ir = ref.search(q1, 10)
    assert torch.equal(I1, Ir) and torch.equal(D1, Dr)
    xh = O.synth(1234, 0, 300_000, 384)          # the same rows on the CPU, for the oracle
    assert_knn_matches(D1.cpu().numpy(), I1.cpu().numpy(), xh, O.synth(5678, 0, 1, 384), 10, what="noalloc nq=1")

    # second call in a row reuses the array with other sizes: more queries (two query groups),
    # larger k; the previous call's lists and ticket must not leak into it
    q2 = b200knn.synth(999, 0, 11, 384, device=dev)
    out2 = (torch.empty((11, 100), dtype=torch.float32, device=x.device),
            torch.empty((11, 100), dtype=torch.int64, device=x.device))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]       # outputs exist already: the caller owns them
    rc, D2, I2 = _noalloc_search(L, dev, x, q2, 100, stream, out=out2)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    assert free_after >= free_before, "the no-alloc entry point allocated device memory"
    Dr, Ir = ref.search(q2, 100)
    assert torch.equal(I2, Ir) and torch.equal(D2, Dr)
    assert_knn_matches(D2.cpu().numpy(), I2.cpu().numpy(), xh, O.synth(999, 0, 11, 384), 100, what="noalloc nq=11")

    # and once more the first query: same answer as the first time
    rc, D3, I3 = _noalloc_search(L, dev, x, q1, 10, stream)
    assert rc == 0 and torch.equal(I3, I1) and torch.equal(D3, D1)

    # k = 1024 is the documented maximum and fits; beyond it the call is refused with a message
    rc, D4, I4 = _noalloc_search(L, dev, x, q1, 1024, stream)
    assert rc == 0, _lib.last_error()
    Dr, Ir = ref.search(q1, 1024)
    assert torch.equal(I4, Ir) and torch.equal(D4, Dr)
    rc, _, _ = _noalloc_search(L, dev, x, q1, 1025, stream)
    assert rc == _lib.ELIMIT and "exceeds" in _lib.last_error()

    # fewer rows than k pads like the regular search; zero rows gives all padding
    rc, D5, I5 = _noalloc_search(L, dev, x[:3].contiguous(), q1, 5, stream)
    assert rc == 0 and I5[0, 3:].tolist() == [-1, -1] and sorted(I5[0, :3].tolist()) == [0, 1, 2]
    rc, D6, I6 = _noalloc_search(L, dev, x[:0].contiguous(), q1, 5, stream)
    assert rc == 0 and bool((I6 == -1).all()) and bool((D6 == np.float32(3.4028235e38)).all())
