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
    dist.barrier()
    if rank == 0:
        print("SHARDED-GPU-OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
