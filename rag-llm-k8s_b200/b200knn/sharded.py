"""Row-sharded IndexFlatL2 across the GPUs of one node, one process per GPU.

Launched under torchrun; every rank makes the same calls (SPMD).  Global rows go to the ranks in
blocks of `block` rows, round-robin (block b lives on rank b % world as its local block
b // world) -- the same layout the single-process multi-GPU handle uses (csrc/multi.cu), pure
arithmetic in both directions, so no id table is kept and adds of any size stay balanced.  Each
rank answers `search` with its shard-local top-k (C ABI, one B200; the kernels write straight into
the packed send record), the per-rank records are exchanged by ONE all-gather, and the merge
kernel (K5, `b2k_merge_shards`) reads the gathered records in place, turning shard-local rows into
global row ids on the way.  The all-gather is a one-shot exchange over peer memory where the GPUs
allow it (`b2k_exchange_push`: every rank stores its record straight into every rank's receive buffer
over NVLink and raises a sequence flag; buffers from torch symmetric memory) and an NCCL
`all_gather_into_tensor` otherwise (`B200KNN_EXCHANGE=nccl` forces it): the payload is a megabyte, so
the step is latency-bound -- a single call at 8 ranks takes 1.28 ms with the peer-store exchange and
1.32 ms with NCCL's; in a back-to-back loop the two are equal.
Ids are global row positions, exactly as a single index would number them, so `llm/rag.py`'s rule "row id == index into the metadata list" (rag.py:81, :119) still holds.

The reference is single-process and has no counterpart of this file; the behaviour it must
reproduce is that of one `faiss.IndexFlatL2` holding all rows (rag.py:61-86, :116).  An index that
must be driven from ONE process (rag.py is one Flask process) uses `IndexFlatL2(d, devices=[...])`
instead; this class is the one-process-per-GPU form the benchmarks launch.
"""

import ctypes as C
import os

import numpy as np
import torch
import torch.distributed as dist

CYCLIC_BLOCK = 4096


def split_range(n, world, rank):
    """Rows [lo, hi) of n rows a rank would keep under a contiguous split (kept for callers that shard that way)."""
    per = -(-n // world) if n else 0
    lo = min(rank * per, n)
    hi = min(lo + per, n)
    return lo, hi


def cyclic_pieces(n0, n, world, rank, block):
    """[(global first row, count)] of the global rows [n0, n0+n) that live on `rank`."""
    out = []
    g = n0
    end = n0 + n
    while g < end:
        e = min((g // block + 1) * block, end)
        if (g // block) % world == rank:
            out.append((g, e - g))
        g = e
    return out


class ShardedIndexFlatL2:
    is_trained = True
    metric_type = 1

    def __init__(self, d, device=None, group=None, block=CYCLIC_BLOCK, _local=None, _merge=None):
        """`_local` / `_merge` replace the CUDA shard index and merge kernel; only the CPU (gloo)
        tests of the host logic use them."""
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.block = int(block)
        self._d = int(d)
        self._cuda = _local is None
        if self._cuda:
            from . import IndexFlatL2, default_device
            self.device = default_device() if device is None else int(device)
            self.local = IndexFlatL2(d, device=self.device)
            self._tdev = torch.device("cuda", self.device)
            self._flag = None
            self._flag_ptr = 0
            self._flags_pin = torch.zeros(self.world + 1, dtype=torch.int64, pin_memory=True)   # shard flags + exchange error
            self._seq = 0
            self._p2p = self._p2p_init() if os.environ.get("B200KNN_EXCHANGE", "p2p") != "nccl" else None
        else:
            self.device = None
            self.local = _local
            self._tdev = torch.device("cpu")
            # a stand-in may bring its own overflow word (int32[2] tensor) to exercise the repeat protocol
            self._flag = getattr(_local, "screen_flag", None)
        self._merge = _merge
        self._bufs = {}          # (nq, k) -> (send, recv): the exchange buffers are reused across searches
        self._tensor_mode = 0
        self._path = 0
        self._ntotal = 0

    # ---- faiss surface -------------------------------------------------------------------
    @property
    def d(self):
        return self._d

    @property
    def ntotal(self):
        return self._ntotal

    def add(self, x):
        """Every rank passes the same batch; each keeps the blocks of the global row order that are its own."""
        n = int(x.shape[0])
        assert x.shape[1] == self._d, "add: dimension mismatch"
        pieces = cyclic_pieces(self._ntotal, n, self.world, self.rank, self.block)
        if pieces:
            parts = [x[g - self._ntotal: g - self._ntotal + c] for g, c in pieces]
            if len(parts) == 1:
                mine = parts[0]
            elif torch.is_tensor(x):
                mine = torch.cat(parts, 0)
            else:
                mine = np.concatenate([np.asarray(p) for p in parts], 0)
            self.local.add(mine)
        self._ntotal += n

    def add_synthetic_global(self, n, seed):
        """Benchmark helper: global rows [ntotal, ntotal+n) get the synthetic stream's rows of the same
        numbers, each rank generating only its own blocks on its own GPU (100M x 768 never exists in one place)."""
        pieces = cyclic_pieces(self._ntotal, n, self.world, self.rank, self.block)
        if pieces:
            self.local.reserve(self.local.ntotal + sum(c for _, c in pieces))
            for g, c in pieces:
                self.local.add_synthetic(c, seed, g)
        self._ntotal += n

    def reset(self):
        self.local.reset()
        self._ntotal = 0

    def set_option(self, key, value):
        if key == "tensor_mode":
            self._tensor_mode = int(value)
        if key == "path":
            self._path = int(value)
        self.local.set_option(key, value)

    def stats(self):
        return self.local.stats()

    def to_global(self, I):
        """Shard-local row ids (as `self.local.search` returns them) -> global row ids (-1 stays -1)."""
        B, W = self.block, self.world
        if torch.is_tensor(I):
            g = (torch.div(I, B, rounding_mode="floor") * W + self.rank) * B + I % B
            return torch.where(I >= 0, g, I)
        I = np.asarray(I)
        return np.where(I >= 0, ((I // B) * W + self.rank) * B + I % B, I)

    def search(self, x, k, D=None, I=None):
        """(D, I) of the k nearest rows over ALL shards; identical on every rank.

        numpy in -> numpy out (host buffers: the queries are copied to the GPU and the results back, as
        `llm/rag.py:116` has it); page-locked arrays -- the queries, and preallocated outputs `D` / `I` as
        in `IndexFlatL2.search` -- are copied asynchronously, so the whole call synchronises once."""
        k = int(k)
        assert k > 0, "search: k must be positive"
        as_numpy = not torch.is_tensor(x)
        xq = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)) if as_numpy else x
        xq = xq.to(self._tdev, torch.float32, non_blocking=True).contiguous()
        assert xq.dim() == 2 and xq.shape[1] == self._d, "search: dimension mismatch"
        host_out = None
        if as_numpy and self._cuda and D is not None and I is not None:
            Dt, It = torch.as_tensor(D), torch.as_tensor(I)
            assert Dt.dtype == torch.float32 and It.dtype == torch.int64 and tuple(Dt.shape) == (xq.shape[0], k) == tuple(It.shape) \
                and Dt.is_contiguous() and It.is_contiguous(), "search: D / I must be C-contiguous (nq, k) float32 / int64 arrays"
            if Dt.is_pinned() and It.is_pinned():
                host_out = (Dt, It)
        Dm, Im, redo = self._search_once(xq, k, host_out)
        # some rank's screening lists could not hold every candidate (masses of near-duplicates at the
        # k-th distance, norms that dwarf the distances): every rank saw the same flags, so every rank
        # repeats -- first with the 3xTF32 product as the screen, then with the exact scan kernel
        for key, val, back in (("tensor_mode", 4, self._tensor_mode), ("path", 1, self._path)):
            if not redo:
                break
            self.local.set_option(key, val)
            try:
                Dm, Im, redo = self._search_once(xq, k, host_out)
            finally:
                self.local.set_option(key, back)
        if host_out is not None:
            return D, I                       # filled before the call's one synchronisation
        if as_numpy:
            Dn, In = torch.as_tensor(Dm).cpu().numpy(), torch.as_tensor(Im).cpu().numpy()
            if D is not None and I is not None:
                D[...] = Dn
                I[...] = In
                return D, I
            return Dn, In
        return Dm, Im

    # ---- exchange over peer memory (torch symmetric memory provides the peer-mapped buffers) -------
    def _all_ok(self, ok):
        t = torch.tensor([1 if ok else 0], device=self._tdev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    def _symm(self, numel, dtype):
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(numel, dtype=dtype, device=self._tdev)
        t.zero_()
        torch.cuda.synchronize(self.device)
        hdl = symm.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
        return t, hdl

    def _p2p_init(self):
        """Flag rows [2][world] (one per buffer half) in symmetric memory; None (-> NCCL) unless every rank succeeds."""
        st = None
        try:
            flags, hdl = self._symm(2 * self.world, torch.int32)
            st = dict(flags=flags, hdl=hdl,
                      flag_ptrs=torch.tensor([int(p) for p in hdl.buffer_ptrs], dtype=torch.int64, device=self._tdev),
                      done=torch.zeros(self.world, dtype=torch.int32, device=self._tdev),
                      err=torch.zeros(2, dtype=torch.int32, device=self._tdev))
        except Exception:      # noqa: BLE001 -- no symmetric memory on this box / build: the NCCL exchange serves
            st = None
        if not self._all_ok(st is not None):
            return None
        dist.barrier(group=self.group)     # every rank's flags are zero before anyone may raise one
        return st

    def _buffers(self, nq, k):
        nk = nq * k
        words = nk + (nk + 1) // 2      # nk int64 ids, then nk float32 distances (8-byte aligned), then one flag word
        slot = (words + 2) & ~1         # record + flag word, padded to 16 bytes (the peer exchange copies 16 bytes at a time)
        b = self._bufs.get((nq, k))
        if b is None:
            # symmetric-memory receive buffers are collective allocations and are never freed: at most 8 shapes get
            # one (every rank makes the same sequence of calls, so every rank decides alike); further shapes, and
            # everything when the peer exchange is off, use ordinary buffers for the NCCL exchange, recycled
            n_symm = sum(1 for v in self._bufs.values() if v[3] is not None)
            if len(self._bufs) - n_symm >= 8:
                for key in [key for key, v in self._bufs.items() if v[3] is None]:
                    del self._bufs[key]
            send = torch.zeros(slot, dtype=torch.int64, device=self._tdev)
            b = None
            if self._cuda and self._p2p is not None and n_symm < 8:
                # receive buffer in symmetric memory: two halves (calls alternate) of `world` slots; a collective
                # allocation, made by every rank at its first search of this shape
                try:
                    recv, hdl = self._symm(2 * self.world * slot, torch.int64)
                    half = self.world * slot * 8
                    dst = [torch.tensor([int(p) + h * half for p in hdl.buffer_ptrs], dtype=torch.int64, device=self._tdev) for h in (0, 1)]
                    ok = True
                except Exception:  # noqa: BLE001
                    ok = False
                if self._all_ok(ok):
                    dist.barrier(group=self.group)
                    b = (send, recv, slot, (hdl, dst))
                else:
                    self._p2p = None
            if b is None:
                b = (send, torch.empty(self.world * slot, dtype=torch.int64, device=self._tdev), slot, None)
            self._bufs[(nq, k)] = b
        return nk, words, b

    def _search_once(self, xq, k, host_out=None):
        nq = xq.shape[0]
        nk, words, (send, recv, slot, p2p) = self._buffers(nq, k)
        I_loc = send[:nk].view(nq, k)
        D_loc = send[nk:words].view(torch.float32)[:nk].view(nq, k)
        if self._cuda:
            return self._search_once_cuda(xq, k, nq, nk, words, send, recv, slot, p2p, D_loc, I_loc, host_out)
        # ---- stand-in path (CPU tests of the host logic): same record layout, ids made global before the exchange
        if self._flag is not None:
            self._flag.zero_()
        D, I = self.local.search(xq, k)
        D_loc.copy_(torch.as_tensor(D))
        I_loc.copy_(torch.as_tensor(self.to_global(torch.as_tensor(I))))
        if self._flag is not None:
            send[words:].copy_(self._flag.view(torch.int64))
        else:
            send[words:].zero_()
        dist.all_gather_into_tensor(recv, send, group=self.group)     # the one collective of the path
        rv = recv.view(self.world, slot)
        Icat = rv[:, :nk].reshape(self.world, nq, k)
        Dcat = rv[:, nk:words].contiguous().view(torch.float32)[:, :nk].reshape(self.world, nq, k)
        Dm, Im = self._merge(Dcat, Icat)
        redo = self._flag is not None and bool((rv[:, words] != 0).any().item())
        return Dm, Im, redo

    def _search_once_cuda(self, xq, k, nq, nk, words, send, recv, slot, p2p, D_loc, I_loc, host_out=None):
        from . import _lib, merge_shards
        # the screening-overflow word of the shard search IS the last word of the send record: with it
        # installed the search does not synchronise, the exchange and the merge are queued behind it
        # while it runs, and the word travels with the all-gather
        fp = send.data_ptr() + 8 * words
        if fp != self._flag_ptr:
            self.local.set_option("screen_flag_ptr", fp)
            self._flag_ptr = fp
        send[words:].zero_()
        self.local.search(xq, k, D=D_loc, I=I_loc)             # kernels write the record in place
        W = self.world
        if p2p is not None:
            # one-shot all-gather: this rank's record goes straight into slot `rank` of every rank's receive
            # buffer (peer stores over NVLink), then the sequence flag; the merge waits for all W flags
            hdl, dst = p2p
            L = _lib.lib()
            self._seq += 1
            seq, half = self._seq, self._seq & 1
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            P = self._p2p
            _lib.check(L.b2k_exchange_push(self.device, C.c_void_p(send.data_ptr()), slot * 8, C.c_void_p(dst[half].data_ptr()),
                                           C.c_void_p(P["flag_ptrs"].data_ptr()), W, self.rank, slot * 8, half * W + self.rank,
                                           seq, C.c_void_p(P["done"].data_ptr()), st))
            _lib.check(L.b2k_exchange_wait(self.device, C.c_void_p(P["flags"].data_ptr() + 4 * half * W), W, seq,
                                           C.c_void_p(P["err"].data_ptr()), st))
            base = recv.data_ptr() + half * W * slot * 8
            rflags = recv[half * W * slot:(half + 1) * W * slot][words::slot]
        else:
            dist.all_gather_into_tensor(recv, send, group=self.group)     # the one collective of the path
            base = recv.data_ptr()
            rflags = recv[words::slot]
        Dm = torch.empty((nq, k), dtype=torch.float32, device=self._tdev)
        Im = torch.empty((nq, k), dtype=torch.int64, device=self._tdev)
        # the merge kernel reads the gathered records where they are and turns shard-local rows into global ids
        merge_shards(base + 8 * nk, 2 * slot, base, slot, W, nq, k, Dm, Im, cyclic_block=self.block)
        self._flags_pin[:W].copy_(rflags, non_blocking=True)
        if p2p is not None:
            self._flags_pin[W:].copy_(self._p2p["err"][:1], non_blocking=True)
        if host_out is not None:              # page-locked outputs: the read-back rides in front of the one synchronisation
            host_out[0].copy_(Dm, non_blocking=True)
            host_out[1].copy_(Im, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        if p2p is not None and int(self._flags_pin[W]) != 0:
            raise RuntimeError("sharded search: rank %d never delivered its record (exchange timed out)" % (int(self._flags_pin[W]) - 1))
        redo = bool(self._flags_pin[:W].any())
        return Dm, Im, redo
