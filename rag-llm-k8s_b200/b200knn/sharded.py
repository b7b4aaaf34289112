"""Row-sharded IndexFlatL2 across the GPUs of one node.

One process per GPU (torchrun); every rank makes the same calls (SPMD).  Each rank keeps a
contiguous slice of every `add` batch in its own `IndexFlatL2` (C ABI, one B200) and answers
`search` with its shard-local top-k; the per-rank (D, I) are exchanged with ONE NCCL all-gather
of a packed buffer and combined by the merge kernel (K5, `b2k_merge_topk`).  Ids are global row
positions, exactly as a single index would number them, so `llm/rag.py`'s rule "row id == index
into the metadata list" (rag.py:81, :119) still holds.

The reference is single-process and has no counterpart of this file; the behaviour it must
reproduce is that of one `faiss.IndexFlatL2` holding all rows (rag.py:61-86, :116).
"""

import numpy as np
import torch
import torch.distributed as dist


def split_range(n, world, rank):
    """Rows [lo, hi) of a batch of n rows kept by `rank`: contiguous blocks of ceil(n/world)."""
    per = -(-n // world) if n else 0
    lo = min(rank * per, n)
    hi = min(lo + per, n)
    return lo, hi


class ShardedIndexFlatL2:
    is_trained = True
    metric_type = 1

    def __init__(self, d, device=None, group=None, _local=None, _merge=None):
        """`_local` / `_merge` replace the CUDA shard index and merge kernel; only the CPU (gloo)
        tests of the host logic use them."""
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self._d = int(d)
        if _local is None:
            from . import IndexFlatL2, default_device
            self.device = default_device() if device is None else int(device)
            self.local = IndexFlatL2(d, device=self.device)
            self._tdev = torch.device("cuda", self.device)
            # screening-overflow word owned by this object (see b2k_set_option "screen_flag_ptr")
            self._flag = torch.zeros(2, dtype=torch.int32, device=self._tdev)
            self.local.set_option("screen_flag_ptr", self._flag.data_ptr())
        else:
            self.device = None
            self.local = _local
            self._tdev = torch.device("cpu")
            # a stand-in may bring its own overflow word (int32[2] tensor) to exercise the repeat protocol
            self._flag = getattr(_local, "screen_flag", None)
        self._merge = _merge
        self._tensor_mode = 0
        self._path = 0
        self._ntotal = 0
        # segments of this rank's shard: local rows [ls, ls+cnt) are global rows [gs, gs+cnt)
        self._seg_local = []
        self._seg_global = []

    # ---- faiss surface -------------------------------------------------------------------
    @property
    def d(self):
        return self._d

    @property
    def ntotal(self):
        return self._ntotal

    def _note_segment(self, gstart, count):
        if count <= 0:
            return
        lstart = self.local.ntotal - count
        if self._seg_local and self._seg_global[-1] + (lstart - self._seg_local[-1]) == gstart:
            return  # contiguous with the previous segment
        self._seg_local.append(lstart)
        self._seg_global.append(gstart)

    def add(self, x):
        """Every rank passes the same batch; rank r keeps rows split_range(n, world, r)."""
        n = int(x.shape[0])
        assert x.shape[1] == self._d, "add: dimension mismatch"
        lo, hi = split_range(n, self.world, self.rank)
        if hi > lo:
            self.local.add(x[lo:hi])
            self._note_segment(self._ntotal + lo, hi - lo)
        self._ntotal += n

    def add_synthetic_global(self, n, seed):
        """Benchmark helper: global rows [ntotal, ntotal+n) of the synthetic stream, each rank
        generating only its own slice on its own GPU (100M x 768 never exists in one place)."""
        lo, hi = split_range(n, self.world, self.rank)
        if hi > lo:
            self.local.reserve(self.local.ntotal + hi - lo)
            self.local.add_synthetic(hi - lo, seed, self._ntotal + lo)
            self._note_segment(self._ntotal + lo, hi - lo)
        self._ntotal += n

    def reset(self):
        self.local.reset()
        self._ntotal = 0
        self._seg_local, self._seg_global = [], []

    def set_option(self, key, value):
        if key == "tensor_mode":
            self._tensor_mode = int(value)
        if key == "path":
            self._path = int(value)
        self.local.set_option(key, value)

    def stats(self):
        return self.local.stats()

    def _to_global(self, I):
        """Shard-local row ids -> global ids (-1 stays -1)."""
        if not self._seg_local:
            return I
        if len(self._seg_local) == 1:
            return torch.where(I >= 0, I + (self._seg_global[0] - self._seg_local[0]), I)
        ls = torch.tensor(self._seg_local, dtype=torch.int64, device=I.device)
        delta = torch.tensor([g - l for g, l in zip(self._seg_global, self._seg_local)], dtype=torch.int64,
                             device=I.device)
        seg = torch.searchsorted(ls, I.clamp(min=0), right=True) - 1
        return torch.where(I >= 0, I + delta[seg], I)

    def to_global(self, I):
        """Shard-local row ids (as `self.local.search` returns them) -> global row ids."""
        return self._to_global(I)

    def search(self, x, k):
        """(D, I) of the k nearest rows over ALL shards; identical on every rank."""
        k = int(k)
        assert k > 0, "search: k must be positive"
        as_numpy = not torch.is_tensor(x)
        xq = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)) if as_numpy else x
        xq = xq.to(self._tdev, torch.float32).contiguous()
        assert xq.dim() == 2 and xq.shape[1] == self._d, "search: dimension mismatch"
        Dm, Im, redo = self._search_once(xq, k)
        # some rank's screening lists could not hold every candidate (masses of near-duplicates at the
        # k-th distance, norms that dwarf the distances): every rank saw the same flags, so every rank
        # repeats -- first with the 3xTF32 product as the screen, then with the exact scan kernel
        for key, val, back in (("tensor_mode", 4, self._tensor_mode), ("path", 1, self._path)):
            if not redo:
                break
            self.local.set_option(key, val)
            try:
                Dm, Im, redo = self._search_once(xq, k)
            finally:
                self.local.set_option(key, back)
        if as_numpy:
            return torch.as_tensor(Dm).cpu().numpy(), torch.as_tensor(Im).cpu().numpy()
        return Dm, Im

    def _search_once(self, xq, k):
        nq = xq.shape[0]
        nk = nq * k
        # packed send buffer: nk int64 ids, nk float32 distances (8-byte aligned), one flag word
        words = nk + (nk + 1) // 2
        send = torch.empty(words + 1, dtype=torch.int64, device=self._tdev)
        I_loc = send[:nk].view(nq, k)
        D_loc = send[nk:words].view(torch.float32)[:nk].view(nq, k)
        if self._flag is not None:
            self._flag.zero_()
        # with the flag word installed the shard search does not synchronise: the exchange and the merge
        # are queued behind it while it runs, and the screening check travels with the all-gather
        D, I = self.local.search(xq, k)
        D_loc.copy_(torch.as_tensor(D))
        I_loc.copy_(self._to_global(torch.as_tensor(I).to(self._tdev)))
        if self._flag is not None:
            send[words:].copy_(self._flag.view(torch.int64))
        else:
            send[words:].zero_()
        recv = torch.empty(self.world * (words + 1), dtype=torch.int64, device=self._tdev)
        dist.all_gather_into_tensor(recv, send, group=self.group)     # the one collective of the path
        recv = recv.view(self.world, words + 1)
        Icat = recv[:, :nk].reshape(self.world, nq, k)
        Dcat = recv[:, nk:words].contiguous().view(torch.float32)[:, :nk].reshape(self.world, nq, k)
        if self._merge is not None:
            Dm, Im = self._merge(Dcat, Icat)
        else:
            from . import merge_topk
            Dm, Im = merge_topk(Dcat, Icat)
        redo = self._flag is not None and bool((recv[:, words] != 0).any().item())
        return Dm, Im, redo
