"""Worker for test_sharded_gloo.py: run under torch.distributed.run with the gloo backend (CPU).

The CUDA shard index and merge kernel are replaced by oracle-backed stand-ins so that the HOST
logic of ShardedIndexFlatL2 (row split, id remapping across several adds, the packed all-gather,
SPMD result equality) is exercised without a GPU.  The product path never takes these stand-ins."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
from oracle import oracle as O  # noqa: E402
from b200knn.sharded import ShardedIndexFlatL2, cyclic_pieces, split_range  # noqa: E402


class OracleShard:
    def __init__(self, d):
        self.x = np.zeros((0, d), np.float32)

    @property
    def ntotal(self):
        return self.x.shape[0]

    def add(self, x):
        self.x = np.concatenate([self.x, np.asarray(x, np.float32)], 0)

    def reserve(self, n):
        pass

    def add_synthetic(self, n, seed, first_row=0):
        self.add(O.synth(seed, first_row, n, self.x.shape[1]))

    def reset(self):
        self.x = self.x[:0]

    def search(self, xq, k):
        D, I = O.search(self.x, np.asarray(xq, np.float32), k)
        return torch.from_numpy(D), torch.from_numpy(I)


class OverflowingShard(OracleShard):
    """Stand-in whose cheap screen 'overflows' on ONE rank: it raises its flag word and returns rubbish,
    as the CUDA shard does when its screening lists cannot hold every candidate.  With tensor_mode 4 it
    overflows again (if told to); with path 1 (the scan) it answers exactly."""

    def __init__(self, d, bad_rank, fail_mode4):
        super().__init__(d)
        self.screen_flag = torch.zeros(2, dtype=torch.int32)
        self.opts = {"tensor_mode": 0, "path": 0}
        self.bad = bad_rank
        self.fail_mode4 = fail_mode4
        self.calls = []

    def set_option(self, key, value):
        self.opts[key] = int(value)

    def search(self, xq, k):
        self.calls.append((self.opts["tensor_mode"], self.opts["path"]))
        exact = self.opts["path"] == 1 or (self.opts["tensor_mode"] == 4 and not self.fail_mode4)
        if self.bad and not exact:
            self.screen_flag[0] = 1
            nq = np.asarray(xq).shape[0]
            return torch.zeros((nq, k)), torch.zeros((nq, k), dtype=torch.int64)
        return super().search(xq, k)


def oracle_merge(Dcat, Icat):
    D, I = O.merge(Dcat.numpy(), Icat.numpy())
    return torch.from_numpy(D), torch.from_numpy(I)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    d = 24
    idx = ShardedIndexFlatL2(d, block=16, _local=OracleShard(d), _merge=oracle_merge)
    x = O.synth(11, 0, 1003, d)
    # several adds of awkward sizes (including fewer rows than ranks) that straddle the 16-row blocks of the cyclic layout
    cuts = [0, 1, 2, 400, 401, 1003]
    for a, b in zip(cuts[:-1], cuts[1:]):
        idx.add(x[a:b])
    assert idx.ntotal == 1003
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([idx.local.ntotal]))
    assert sum(int(c) for c in counts) == 1003
    for nq, k in ((1, 5), (7, 10), (25, 3), (2, 1003 + 5)):
        q = O.synth(12, 0, nq, d)
        D, I = idx.search(q, k)
        Dr, Ir = O.search(x, q, k)
        assert np.array_equal(I, Ir), (rank, nq, k, I[0, :8], Ir[0, :8])
        assert np.array_equal(D, Dr), (rank, nq, k)
    # the benchmark's way of filling the shards
    idx.reset()
    idx.add_synthetic_global(777, 5)
    idx.add_synthetic_global(100, 5)          # second call continues the id space, not the stream
    xa = np.concatenate([O.synth(5, 0, 777, d), O.synth(5, 777, 100, d)], 0)
    q = O.synth(13, 0, 3, d)
    D, I = idx.search(q, 9)
    Dr, Ir = O.search(xa, q, 9)
    assert np.array_equal(I, Ir) and np.array_equal(D, Dr)
    # empty index pads
    idx.reset()
    D, I = idx.search(q, 4)
    assert np.all(I == -1) and np.all(D == O.FLT_MAX)
    assert split_range(10, 4, 3) == (9, 10) and split_range(2, 4, 3) == (2, 2)
    assert cyclic_pieces(10, 30, 2, 1, 16) == [(16, 16)] and cyclic_pieces(10, 30, 2, 0, 16) == [(10, 6), (32, 8)]
    ids = np.arange(40)
    mine = np.concatenate([np.arange(g, g + c) for g, c in cyclic_pieces(0, 40, world, rank, 16)] or [np.zeros(0, np.int64)])
    assert np.array_equal(idx.to_global(np.arange(len(mine))), mine)     # local row -> global row is the inverse of the split
    # screening overflow on the last rank only: its flag travels with the all-gather, EVERY rank repeats
    # (3xTF32 screen, then the scan), the options are restored, the answer is the single index's
    for fail_mode4, expect in ((False, [(0, 0), (4, 0)]), (True, [(0, 0), (4, 0), (0, 1)])):
        shard = OverflowingShard(d, bad_rank=(rank == world - 1), fail_mode4=fail_mode4)
        idx2 = ShardedIndexFlatL2(d, block=16, _local=shard, _merge=oracle_merge)
        idx2.add(x)
        q = O.synth(14, 0, 6, d)
        D, I = idx2.search(q, 7)
        Dr, Ir = O.search(x, q, 7)
        assert np.array_equal(I, Ir) and np.array_equal(D, Dr), (rank, fail_mode4)
        assert shard.calls == expect, (rank, shard.calls)
        assert shard.opts == {"tensor_mode": 0, "path": 0}
    dist.barrier()
    if rank == 0:
        print("SHARDED-OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
