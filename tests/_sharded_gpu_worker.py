"""Worker for test_gpu_sharded.py: torchrun, NCCL backend, one rank per GPU; the real CUDA path."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as O  # noqa: E402
from util import assert_knn_matches  # noqa: E402
from b200knn.sharded import ShardedIndexFlatL2  # noqa: E402


def main():
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    rank, world = dist.get_rank(), dist.get_world_size()
    d = 128
    idx = ShardedIndexFlatL2(d, device=lr)
    x = O.synth(21, 0, 60011, d)
    for a, b in ((0, 1), (1, 30000), (30000, 60011)):
        idx.add(x[a:b])
    for nq, k, expanded in ((1, 10, False), (9, 5, False), (64, 100, True)):
        q = O.synth(22, 0, nq, d)
        D, I = idx.search(q, k)                      # numpy in, numpy out
        assert_knn_matches(D, I, x, q, k, expanded=expanded, what="sharded nq=%d rank=%d" % (nq, rank))
        Dt, It = idx.search(torch.from_numpy(q).cuda(), k)   # device tensors
        assert np.array_equal(It.cpu().numpy(), I) and np.array_equal(Dt.cpu().numpy(), D)
    idx.reset()
    idx.add_synthetic_global(200000, 1234)
    xs = O.synth(1234, 0, 200000, d)
    q = O.synth(5678, 0, 40, d)
    D, I = idx.search(q, 10)
    assert_knn_matches(D, I, xs, q, 10, expanded=True, what="sharded synthetic")
    # screening overflow on ONE shard only (the near-duplicates are the rows the last rank keeps): the flag
    # travels with the all-gather, every rank repeats the search (3xTF32 screen, then the scan), results stay exact
    idx.reset()
    rng = np.random.default_rng(0)
    base = rng.standard_normal((1, 64)).astype(np.float32)
    per = 20000
    z = rng.standard_normal((per * world, 64)).astype(np.float32)
    idx2 = ShardedIndexFlatL2(64, device=lr)
    last = np.where((np.arange(per * world) // idx2.block) % world == world - 1)[0]   # the rows the last rank keeps
    z[last] = base + rng.standard_normal((len(last), 64)).astype(np.float32) * 1e-6
    z[last[::1000]] = rng.standard_normal((len(last[::1000]), 64)).astype(np.float32)
    idx2.add(z)
    qz = (base + rng.standard_normal((30, 64)).astype(np.float32) * 0.05).astype(np.float32)
    D, I = idx2.search(qz, 10)
    assert_knn_matches(D, I, z, qz, 10, expanded=True, what="sharded overflow redo rank=%d" % rank)
    st = idx2.stats()
    assert st["path"] == "scan" or st.get("tensor_mode") == "3xtf32screen+rerank", st
    idx.add_synthetic_global(200000, 1234)               # the first index again: screening mode, no redo
    D2, I2 = idx.search(q, 10)
    assert idx.stats()["tensor_mode"] == "bf16screen+rerank"
    assert_knn_matches(D2, I2, xs, q, 10, expanded=True, what="sharded synthetic again")
    dist.barrier()
    # numpy in, page-locked outputs: filled in place, one synchronisation
    Dp = torch.empty((40, 10), dtype=torch.float32, pin_memory=True).numpy()
    Ip = torch.empty((40, 10), dtype=torch.int64, pin_memory=True).numpy()
    D3, I3 = idx.search(q, 10, D=Dp, I=Ip)
    assert D3 is Dp and I3 is Ip and np.array_equal(Ip, I2) and np.array_equal(Dp, D2)
    dist.barrier()
    if rank == 0:
        print("SHARDED-GPU-OK world=%d exchange=%s" % (world, "p2p" if idx._p2p is not None else "nccl"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
