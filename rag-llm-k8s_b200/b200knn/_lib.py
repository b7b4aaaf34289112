"""ctypes binding of libb200knn.so (the C ABI in include/b200knn.h).

The library is built in-tree by `rag-llm-k8s_b200/build.py`; importing this module never
compiles anything and never falls back to a CPU implementation: if the shared object is
missing, or no B200 is usable, the first call raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200knn.so")

OK, EINVAL, ECUDA, EIO, ENOMEM, ELIMIT = 0, -1, -2, -3, -4, -5
MAX_K = 1024
CYCLIC_BLOCK = 4096   # B2K_CYCLIC_BLOCK

_f32p = C.POINTER(C.c_float)
_i64p = C.POINTER(C.c_int64)

# name -> (restype, argtypes); must list every symbol include/b200knn.h declares
SIGNATURES = {
    "b2k_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "b2k_destroy": (C.c_int, [C.c_void_p]),
    "b2k_add": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]),
    "b2k_search": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "b2k_ntotal": (C.c_int64, [C.c_void_p]),
    "b2k_d": (C.c_int, [C.c_void_p]),
    "b2k_reset": (C.c_int, [C.c_void_p]),
    "b2k_write": (C.c_int, [C.c_void_p, C.c_char_p]),
    "b2k_write_append": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "b2k_read": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]),
    "b2k_create_multi": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    "b2k_read_multi": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    "b2k_view": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]),
    "b2k_last_error": (C.c_char_p, []),
    "b2k_reserve": (C.c_int, [C.c_void_p, C.c_int64]),
    "b2k_get_rows": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int]),
    "b2k_synth": (C.c_int, [C.c_int, C.c_uint64, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]),
    "b2k_add_synthetic": (C.c_int, [C.c_void_p, C.c_int64, C.c_uint64, C.c_int64]),
    "b2k_merge_topk": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "b2k_merge_shards": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                   C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2k_exchange_push": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_int,
                                    C.c_uint32, C.c_void_p, C.c_void_p]),
    "b2k_exchange_wait": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p]),
    "b2k_search_rows_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2k_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "b2k_get_stat": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]),
    "b2k_launch_count": (C.c_int64, []),
    "b2k_version": (C.c_int, []),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libb200knn.so is not built (%s missing): run `python rag-llm-k8s_b200/build.py`; "
                "b200knn has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error():
    msg = lib().b2k_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc):
    """Map a C return code to the exception faiss's wrapper would raise."""
    if rc == OK:
        return
    msg = last_error() or ("libb200knn error %d" % rc)
    if rc == EINVAL:
        raise AssertionError(msg)
    if rc == ENOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)
