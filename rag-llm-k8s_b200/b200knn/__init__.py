"""b200knn -- drop-in for the `faiss` names used by oscka/rag-llm-k8s `llm/rag.py`.

    import b200knn as faiss                      # the one edit to llm/rag.py (line 6)
    index = faiss.IndexFlatL2(d)                 # rag.py:61, :71
    index.add(vectors_np)                        # rag.py:80
    D, I = index.search(q.reshape(1, -1), 5)     # rag.py:116
    index.ntotal, index.d                        # rag.py:86, :190-191
    faiss.write_index(index, path)               # rag.py:62, :82
    index = faiss.read_index(path)               # rag.py:74, :153, :186

Every call goes through the C ABI of libb200knn.so (include/b200knn.h) onto a B200; there
is no CPU path.  numpy inputs are host buffers (results are fresh numpy arrays, as faiss
returns); torch CUDA tensors are used in place (results are torch tensors on that device,
enqueued on torch's current stream).
"""
import collections
import ctypes as C
import os
import threading

import numpy as np

from . import _lib
from ._lib import MAX_K

__all__ = ["IndexFlatL2", "write_index", "read_index", "METRIC_L2", "default_device", "default_devices",
           "launch_count", "clear_index_cache"]

METRIC_L2 = 1
__version__ = "0.1.0"


def default_device():
    """GPU used by new indexes: $B200KNN_DEVICE, else $LOCAL_RANK (torchrun), else 0."""
    for var in ("B200KNN_DEVICE", "LOCAL_RANK"):
        v = os.environ.get(var)
        if v not in (None, ""):
            return int(v)
    return 0


def default_devices():
    """GPUs a new index spans: $B200KNN_DEVICES ("0,1,2,3": one index row-sharded over those GPUs, driven
    from this one process -- what a single Flask process like llm/rag.py needs to use a whole node), else
    the one GPU of default_device()."""
    v = os.environ.get("B200KNN_DEVICES")
    if v not in (None, ""):
        devs = [int(t) for t in v.replace(" ", "").split(",") if t != ""]
        if devs:
            return devs
    return [default_device()]


def launch_count():
    return int(_lib.lib().b2k_launch_count())


def _is_torch(x):
    return type(x).__module__.split(".")[0] == "torch"


def _as_host_f32(x, d, what):
    x = np.ascontiguousarray(x, dtype=np.float32)
    assert x.ndim == 2, "%s: expected a 2-D array, got shape %s" % (what, x.shape)
    assert x.shape[1] == d, "%s: dimension mismatch, index has d=%d, got %d" % (what, d, x.shape[1])
    return x


def _as_dev_f32(x, d, device, what):
    import torch
    assert x.is_cuda and x.device.index == device, "%s: tensor must live on cuda:%d" % (what, device)
    assert x.dim() == 2 and x.shape[1] == d, "%s: expected shape (n, %d), got %s" % (what, d, tuple(x.shape))
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.to(torch.float32).contiguous()
    return x


def _int_array(devs):
    return (C.c_int * len(devs))(*devs)


class IndexFlatL2:
    """Exact squared-L2 index on one B200 -- or row-sharded over several (`devices=[...]`, or
    $B200KNN_DEVICES), still one object in one process; the surface of `faiss.IndexFlatL2` that rag.py uses."""

    is_trained = True
    metric_type = METRIC_L2
    # provenance for the incremental write_index (see read_index below):
    _origin = None   # (realpath, stat key, rows in that file) while those rows are a prefix of this index

    def __init__(self, d, device=None, devices=None, _handle=None):
        self._h = None
        if devices is None:
            devices = default_devices() if device is None else [int(device)]
        self.devices = [int(v) for v in devices]
        assert self.devices, "IndexFlatL2: empty device list"
        self.device = self.devices[0]          # where device-side inputs / outputs live
        if _handle is not None:
            self._h = _handle
            return
        h = C.c_void_p()
        if len(self.devices) > 1:
            _lib.check(_lib.lib().b2k_create_multi(int(d), _int_array(self.devices), len(self.devices), C.byref(h)))
        else:
            _lib.check(_lib.lib().b2k_create(int(d), self.device, C.byref(h)))
        self._h = h

    def _view(self):
        """An independent index over the rows this one holds now, sharing their storage in HBM (b2k_view)."""
        h = C.c_void_p()
        _lib.check(_lib.lib().b2k_view(self._h, -1, C.byref(h)))
        v = IndexFlatL2(0, devices=self.devices, _handle=h)
        v._origin = self._origin
        return v

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().b2k_destroy(h)
            except Exception:
                pass

    @property
    def d(self):
        return int(_lib.lib().b2k_d(self._h))

    @property
    def ntotal(self):
        return int(_lib.lib().b2k_ntotal(self._h))

    def _stream(self):
        import torch
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def add(self, x):
        """Append rows (rag.py:80).  Ids are row positions, sequential across adds."""
        L = _lib.lib()
        if _is_torch(x) and x.is_cuda:
            x = _as_dev_f32(x, self.d, self.device, "add")
            _lib.check(L.b2k_add(self._h, x.shape[0], C.c_void_p(x.data_ptr()), 1, self._stream()))
            return
        if _is_torch(x):
            x = x.detach().cpu().numpy()
        x = _as_host_f32(x, self.d, "add")
        _lib.check(L.b2k_add(self._h, x.shape[0], x.ctypes.data_as(C.c_void_p), 0, None))

    def search(self, x, k, params=None, D=None, I=None):
        """(D, I) = k nearest stored rows of each query row (rag.py:116).

        Same optional arguments as faiss: `D` / `I` are preallocated outputs (float32 / int64,
        C-contiguous, shape (nq, k)) that are filled and returned; page-locked host arrays are
        copied to / from directly, without the library's staging buffer.  `params` must be None."""
        L = _lib.lib()
        k = int(k)
        assert k > 0, "search: k must be positive"
        assert params is None, "search: IndexFlatL2 takes no search parameters"
        if _is_torch(x) and x.is_cuda:
            import torch
            x = _as_dev_f32(x, self.d, self.device, "search")
            nq = x.shape[0]
            if D is None:
                D = torch.empty((nq, k), dtype=torch.float32, device=x.device)
            if I is None:
                I = torch.empty((nq, k), dtype=torch.int64, device=x.device)
            assert _is_torch(D) and _is_torch(I) and D.is_cuda and I.is_cuda and D.device == x.device and I.device == x.device, \
                "search: D / I must be CUDA tensors on the queries' device"
            assert D.dtype == torch.float32 and I.dtype == torch.int64 and D.is_contiguous() and I.is_contiguous() \
                and tuple(D.shape) == (nq, k) and tuple(I.shape) == (nq, k), "search: D / I must be contiguous (nq, k) float32 / int64"
            _lib.check(L.b2k_search(self._h, nq, C.c_void_p(x.data_ptr()), k, C.c_void_p(D.data_ptr()),
                                    C.c_void_p(I.data_ptr()), 1, self._stream()))
            return D, I
        if _is_torch(x):
            x = x.detach().cpu().numpy()
        x = _as_host_f32(x, self.d, "search")
        nq = x.shape[0]
        if D is None:
            D = np.empty((nq, k), dtype=np.float32)
        if I is None:
            I = np.empty((nq, k), dtype=np.int64)
        assert isinstance(D, np.ndarray) and isinstance(I, np.ndarray) and D.dtype == np.float32 and I.dtype == np.int64 \
            and D.flags["C_CONTIGUOUS"] and I.flags["C_CONTIGUOUS"] and D.shape == (nq, k) and I.shape == (nq, k), \
            "search: D / I must be C-contiguous (nq, k) float32 / int64 arrays"
        _lib.check(L.b2k_search(self._h, nq, x.ctypes.data_as(C.c_void_p), k, D.ctypes.data_as(C.c_void_p),
                                I.ctypes.data_as(C.c_void_p), 0, None))
        return D, I

    def reset(self):
        self._origin = None
        _lib.check(_lib.lib().b2k_reset(self._h))

    def reserve(self, nrows):
        _lib.check(_lib.lib().b2k_reserve(self._h, int(nrows)))

    def prepare(self):
        """Build the bf16 copies the search kernels screen on (+50 % of the row bytes) now rather than inside the
        first search that would profit from them; later adds keep them current."""
        self.set_option("prepare_copies", 1)

    def reconstruct_n(self, i0=0, n=-1):
        if n < 0:
            n = self.ntotal - i0
        out = np.empty((n, self.d), dtype=np.float32)
        _lib.check(_lib.lib().b2k_get_rows(self._h, int(i0), int(n), out.ctypes.data_as(C.c_void_p), 0))
        return out

    def add_synthetic(self, n, seed, first_row=0):
        """Append n rows of the counter-based synthetic stream, generated on the GPU."""
        _lib.check(_lib.lib().b2k_add_synthetic(self._h, int(n), int(seed), int(first_row)))

    def stats(self):
        """{'path': 'scan'|'tensor'|None, 'main_kernel_ms': float|None} for the last search."""
        L = _lib.lib()
        v = C.c_double()
        out = {"path": None, "main_kernel_ms": None}
        if L.b2k_get_stat(self._h, b"last_path", C.byref(v)) == 0:
            out["path"] = {1: "scan", 2: "tensor"}.get(int(v.value))
        if out["path"] == "tensor" and L.b2k_get_stat(self._h, b"tensor_kernel_ms", C.byref(v)) == 0:
            out["main_kernel_ms"] = v.value
        if L.b2k_get_stat(self._h, b"screen_fallbacks", C.byref(v)) == 0:
            out["screen_fallbacks"] = int(v.value)
        if out["path"] == "tensor" and L.b2k_get_stat(self._h, b"tensor_mode_used", C.byref(v)) == 0:
            out["tensor_mode"] = {1: "3xtf32", 2: "screen+rerank", 3: "bf16screen+rerank", 4: "3xtf32screen+rerank"}.get(int(v.value))
            L.b2k_get_stat(self._h, b"screen_fallbacks", C.byref(v))
            out["screen_fallbacks"] = int(v.value)
        if out["path"] == "scan" and L.b2k_get_stat(self._h, b"scan_kernel_ms", C.byref(v)) == 0:
            out["main_kernel_ms"] = v.value
            L.b2k_get_stat(self._h, b"scan_launches", C.byref(v))
            out["launches"] = int(v.value)
        if out["path"] == "scan" and L.b2k_get_stat(self._h, b"scan_rows_used", C.byref(v)) == 0:
            out["scan_rows"] = {1: "fp32", 2: "bf16+rerank"}.get(int(v.value))
        if L.b2k_get_stat(self._h, b"devices", C.byref(v)) == 0:
            out["devices"] = int(v.value)
        return out

    def set_option(self, key, value):
        _lib.check(_lib.lib().b2k_set_option(self._h, key.encode(), int(value)))


# ---- persistence ---------------------------------------------------
# llm/rag.py re-reads the whole index file on EVERY request (rag.py:153, :186) and rewrites the
# whole file on every upload (rag.py:74-84).  With the index in HBM that would put a file read
# plus an H2D copy of the full index in front of every search, so:
#   * read_index keeps the last few loaded files resident, keyed on (realpath, mtime_ns, size,
#     inode); for an unchanged file it returns a fresh VIEW of the resident rows (b2k_view) without
#     touching the disk or the bus.  As with faiss, every call returns its own independent index:
#     rows added through one are invisible to every other (a view appends in place while it is the
#     only one at the end of the shared rows, and takes a private copy of its prefix otherwise), so
#     concurrent requests of a threaded Flask server never see each other's half-finished updates;
#   * write_index of an index that was loaded from `path` and only grew since appends the new
#     rows and rewrites the 45-byte header instead of rewriting everything (the file stays
#     byte-identical to a full write).
# B200KNN_INDEX_CACHE=0 turns the residency off (every read loads the file); =N keeps N files (default 4).
_cache = collections.OrderedDict()
_cache_lock = threading.Lock()


def _cache_max():
    try:
        return max(int(os.environ.get("B200KNN_INDEX_CACHE", "4")), 0)
    except ValueError:
        return 4


def _stat_key(path):
    st = os.stat(path)
    return (st.st_mtime_ns, st.st_size, st.st_ino)


def _remember(rp, key, index):
    """`index` is, right now, exactly the file (rp, key): keep a private snapshot view of it resident."""
    index._origin = (rp, key, index.ntotal)
    n = _cache_max()
    snap = index._view() if n else None      # nobody else holds this handle, so it can never change
    with _cache_lock:
        _cache.pop(rp, None)
        if n:
            _cache[rp] = (key, snap)
            while len(_cache) > n:
                _cache.popitem(last=False)


def clear_index_cache():
    with _cache_lock:
        _cache.clear()


def write_index(index, path):
    """faiss.write_index (rag.py:62, :82): byte-exact IxF2 file.  Full writes go through
    path.tmp + rename; an index that only grew since it was loaded from `path` is appended."""
    L = _lib.lib()
    rp = os.path.realpath(path)
    o = index._origin
    appended = False
    if o is not None and o[0] == rp and o[2] <= index.ntotal and _cache_max():
        try:
            same = _stat_key(rp) == o[1]
        except OSError:
            same = False
        if same:
            appended = L.b2k_write_append(index._h, os.fsencode(path), o[2]) == _lib.OK
    if not appended:
        _lib.check(L.b2k_write(index._h, os.fsencode(path)))
    _remember(rp, _stat_key(rp), index)


def read_index(path, device=None, devices=None):
    """faiss.read_index (rag.py:74, :153, :186); a missing file raises RuntimeError."""
    if devices is None:
        devices = default_devices() if device is None else [int(device)]
    devices = [int(v) for v in devices]
    rp = os.path.realpath(path)
    key = None
    try:
        key = _stat_key(rp)
    except OSError:
        pass                                   # let the library produce faiss's error message
    if key is not None and _cache_max():
        with _cache_lock:
            ent = _cache.get(rp)
            if ent is not None and ent[0] == key and ent[1].devices == devices:
                _cache.move_to_end(rp)
                return ent[1]._view()
    h = C.c_void_p()
    if len(devices) > 1:
        _lib.check(_lib.lib().b2k_read_multi(os.fsencode(path), _int_array(devices), len(devices), C.byref(h)))
    else:
        _lib.check(_lib.lib().b2k_read(os.fsencode(path), devices[0], C.byref(h)))
    idx = IndexFlatL2(0, devices=devices, _handle=h)
    if key is not None and _stat_key(rp) == key:
        _remember(rp, key, idx)
    return idx


def synth(seed, row0, nrows, d, device=None):
    """Synthetic rows as a torch CUDA tensor (same stream as oracle.synth on the CPU)."""
    import torch
    dev = default_device() if device is None else int(device)
    out = torch.empty((nrows, d), dtype=torch.float32, device="cuda:%d" % dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(_lib.lib().b2k_synth(dev, int(seed), int(row0), int(nrows), int(d), C.c_void_p(out.data_ptr()), st))
    return out


def merge_topk(Dl, Il, id_offset=None, device=None):
    """Combine [nlists, nq, k] result sets into the global (D, I) -- the sharded search's last step."""
    L = _lib.lib()
    off = None
    if id_offset is not None:
        off_arr = np.ascontiguousarray(id_offset, dtype=np.int64)
        off = off_arr.ctypes.data_as(C.c_void_p)
    if _is_torch(Dl):
        import torch
        dev = Dl.device.index
        Dl = Dl.contiguous()
        Il = Il.contiguous()
        nl, nq, k = Dl.shape
        D = torch.empty((nq, k), dtype=torch.float32, device=Dl.device)
        I = torch.empty((nq, k), dtype=torch.int64, device=Dl.device)
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(L.b2k_merge_topk(dev, nl, nq, k, C.c_void_p(Dl.data_ptr()), C.c_void_p(Il.data_ptr()), off,
                                    C.c_void_p(D.data_ptr()), C.c_void_p(I.data_ptr()), 1, st))
        return D, I
    dev = default_device() if device is None else int(device)
    Dl = np.ascontiguousarray(Dl, dtype=np.float32)
    Il = np.ascontiguousarray(Il, dtype=np.int64)
    nl, nq, k = Dl.shape
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    _lib.check(L.b2k_merge_topk(dev, nl, nq, k, Dl.ctypes.data_as(C.c_void_p), Il.ctypes.data_as(C.c_void_p), off,
                                D.ctypes.data_as(C.c_void_p), I.ctypes.data_as(C.c_void_p), 0, None))
    return D, I


def merge_shards(Dpk, stride_d, Ipk, stride_i, nlists, nq, k, D, I, cyclic_block=_lib.CYCLIC_BLOCK):
    """Combine step over device-side lists with their own strides (see b2k_merge_shards); torch CUDA tensors."""
    import torch
    dev = D.device.index
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(_lib.lib().b2k_merge_shards(dev, int(nlists), int(nq), int(k), C.c_void_p(Dpk), int(stride_d), C.c_void_p(Ipk),
                                           int(stride_i), int(cyclic_block), C.c_void_p(D.data_ptr()), C.c_void_p(I.data_ptr()), st))
    return D, I
