"""One IndexFlatL2 row-sharded over several GPUs, driven from ONE process (b2k_create_multi behind
`IndexFlatL2(d, devices=[...])` / $B200KNN_DEVICES): llm/rag.py is a single Flask process
(rag.py:204), so this is the form in which its index can use more than one B200.  The behaviour to
match is that of one index holding all rows -- same ids, same (distance, id) order, same IxF2 file.

Device lists: [0, 1] / [0, 1, 2] need that many GPUs; [0, 0] and [0, 0, 0] put several shards on the
same GPU through the library's test hook (B200KNN_TEST_DUP_DEVICES=1), so that a one-GPU box still
runs the whole sharded code path (block-cyclic split, worker threads, packed result buffer, merge
kernel with id translation, persistence in global row order, views)."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from util import assert_knn_matches
from test_gpu_parity import rag_py_call_pattern

pytestmark = pytest.mark.gpu

DEVICE_LISTS = [[0, 0], [0, 0, 0], [0, 1], [0, 1, 2]]


@pytest.fixture(scope="module")
def faiss():
    os.environ["B200KNN_TEST_DUP_DEVICES"] = "1"
    import b200knn
    return b200knn


def _need(devs):
    import torch
    if max(devs) >= torch.cuda.device_count():
        pytest.skip("needs %d GPUs" % (max(devs) + 1))


@pytest.mark.parametrize("devs", DEVICE_LISTS, ids=lambda v: "gpus" + "".join(map(str, v)))
def test_multi_matches_the_oracle(faiss, devs, tmp_path):
    _need(devs)
    d = 96
    n = 3 * 4096 * len(devs) + 1234            # whole cycles of blocks plus a ragged tail
    x = O.synth(61, 0, n, d)
    idx = faiss.IndexFlatL2(d, devices=devs)
    assert idx.ntotal == 0 and idx.d == d
    q1 = O.synth(62, 0, 1, d)
    D, I = idx.search(q1, 4)                   # empty index pads
    assert np.all(I == -1) and np.all(D == O.FLT_MAX)
    cuts = [0, 1, 3, 4095, 4097, 20000, n]     # adds that start and end inside blocks
    for a, b in zip(cuts[:-1], cuts[1:]):
        idx.add(x[a:b])
        assert idx.ntotal == b
    assert np.array_equal(idx.reconstruct_n(4090, 12), x[4090:4102])      # rows come back in global order
    for nq, k in ((1, 5), (9, 10), (64, 100), (300, 7)):
        q = O.synth(62, 0, nq, d)
        D, I = idx.search(q, k)
        assert_knn_matches(D, I, x, q, k, expanded=nq >= 20, what="multi %s nq=%d" % (devs, nq))
    # fewer rows than k over all shards: padding after the merge
    small = faiss.IndexFlatL2(d, devices=devs)
    small.add(x[:3])
    D, I = small.search(q1, 5)
    assert_knn_matches(D, I, x[:3], q1, 5, what="multi pad")
    # without direct peer stores (the path for GPUs without P2P access): same answer
    q = O.synth(62, 0, 64, d)
    D0, I0 = idx.search(q, 10)
    idx.set_option("multi_peer_writes", 0)
    D1, I1 = idx.search(q, 10)
    idx.set_option("multi_peer_writes", 1)
    assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    # persistence: ONE IxF2 file in global row order, byte-identical to the independent writer's, loadable on any device count
    p = str(tmp_path / "multi.index")
    faiss.clear_index_cache()
    faiss.write_index(idx, p)
    ref = str(tmp_path / "ref.index")
    O.write_ixf2(ref, x)
    assert open(p, "rb").read() == open(ref, "rb").read()
    faiss.clear_index_cache()
    again = faiss.read_index(p, devices=devs)
    assert again.ntotal == n
    D2, I2 = again.search(q, 10)
    assert np.array_equal(I0, I2) and np.array_equal(D0, D2)
    one = faiss.read_index(p, devices=[0])
    D3, I3 = one.search(q, 10)
    assert np.array_equal(I0, I3) and np.array_equal(D0, D3)
    # incremental write_index (rag.py:74-84) through the sharded handle
    extra = O.synth(63, 0, 5000, d)
    again.add(extra)
    faiss.write_index(again, p)
    O.write_ixf2(ref, np.concatenate([x, extra]))
    assert open(p, "rb").read() == open(ref, "rb").read()
    # reset
    again.reset()
    assert again.ntotal == 0
    again.add(x[:100])
    D, I = again.search(q1, 3)
    assert_knn_matches(D, I, x[:100], q1, 3, what="multi after reset")


@pytest.mark.parametrize("devs", [[0, 0], [0, 1]], ids=lambda v: "gpus" + "".join(map(str, v)))
def test_multi_device_tensors_views_and_redo(faiss, devs):
    _need(devs)
    import torch
    d, n = 64, 70000
    x = faiss.synth(1234, 0, n, d, device=0)
    idx = faiss.IndexFlatL2(d, devices=devs)
    idx.add(x)                                      # rows live on devices[0]; the other shards fetch their blocks
    xh = O.synth(1234, 0, n, d)
    q = faiss.synth(5678, 0, 40, d, device=0)
    D, I = idx.search(q, 10)
    assert D.is_cuda and I.is_cuda
    assert_knn_matches(D.cpu().numpy(), I.cpu().numpy(), xh, q.cpu().numpy(), 10, expanded=True, what="multi device tensors")
    # a view is an independent index over the same rows (what read_index hands out per request)
    v = idx._view()
    idx.add(O.synth(77, 0, 9000, d))
    assert v.ntotal == n and idx.ntotal == n + 9000
    Dv, Iv = v.search(q, 10)
    assert torch.equal(Iv, I) and torch.equal(Dv, D)
    v.add(O.synth(78, 0, 10, d))                    # the shared rows have moved on: private copy
    assert v.ntotal == n + 10
    assert np.array_equal(v.reconstruct_n(n - 5, 15), np.concatenate([xh[n - 5:], O.synth(78, 0, 10, d)]))
    # screening overflow on some shard: all shards repeat down the exact chain, results stay exact
    rng = np.random.default_rng(0)
    base = rng.standard_normal((1, d)).astype(np.float32)
    z = rng.standard_normal((60000, d)).astype(np.float32)
    z[4096:4096 + 20000] = base + rng.standard_normal((20000, d)).astype(np.float32) * 1e-6
    z[4096:4096 + 20000:1000] = rng.standard_normal((20, d)).astype(np.float32)
    dup = faiss.IndexFlatL2(d, devices=devs)
    dup.add(z)
    qz = (base + rng.standard_normal((30, d)).astype(np.float32) * 0.05).astype(np.float32)
    D, I = dup.search(qz, 10)
    assert_knn_matches(D, I, z, qz, 10, expanded=True, what="multi overflow redo")
    assert dup.stats()["screen_fallbacks"] >= 1


@pytest.mark.parametrize("devs", [[0, 0], [0, 1]], ids=lambda v: "gpus" + "".join(map(str, v)))
def test_rag_py_call_pattern_on_a_sharded_index(faiss, devs, tmp_path, monkeypatch):
    """The one-line edit of INTEGRATION.md plus B200KNN_DEVICES: rag.py's own call sequence, unchanged."""
    _need(devs)
    monkeypatch.setenv("B200KNN_DEVICES", ",".join(map(str, devs)))
    faiss.clear_index_cache()
    assert faiss.default_devices() == devs
    rag_py_call_pattern(faiss, tmp_path)
    idx = faiss.read_index(str(tmp_path / "faiss_index"))
    assert idx.devices == devs and idx.stats().get("devices") == len(devs)


def test_ndev_one_is_the_plain_handle(faiss):
    idx = faiss.IndexFlatL2(16, devices=[0])
    x = O.synth(5, 0, 500, 16)
    idx.add(x)
    q = O.synth(6, 0, 2, 16)
    D, I = idx.search(q, 3)
    assert_knn_matches(D, I, x, q, 3, what="ndev=1")
    assert idx.stats().get("devices") == 1
