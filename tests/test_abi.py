"""CPU checks of the boundary: the C-ABI library loads, exports every symbol the header
declares, and -- when no GPU is present -- fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200knn.h")).read()
    return sorted(set(re.findall(r"B2K_API[^;(]*?\b(b2k_\w+)\s*\(", src)))


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    from b200knn import _lib
    L = _lib.lib()
    names = _header_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), "libb200knn.so does not export %s" % n
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and include/b200knn.h disagree"
    assert L.b2k_version() >= 100


def test_product_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "rag-llm-k8s_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.sep + "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f


@pytest.mark.skipif(_have_gpu(), reason="only meaningful on a box without a GPU")
def test_no_gpu_means_loud_failure_not_fallback():
    import b200knn
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        b200knn.IndexFlatL2(8)
    with pytest.raises(RuntimeError):
        b200knn.merge_topk(np.zeros((1, 1, 1), np.float32), np.zeros((1, 1, 1), np.int64))


def test_argument_errors_map_like_faiss_wrapper():
    from b200knn import _lib
    L = _lib.lib()
    h = C.c_void_p()
    assert L.b2k_create(0, 0, C.byref(h)) == _lib.EINVAL       # d <= 0
    with pytest.raises(AssertionError):
        _lib.check(_lib.EINVAL)
    with pytest.raises(RuntimeError):
        _lib.check(_lib.EIO)
    assert L.b2k_search(None, 1, None, 5, None, None, 0, None) == _lib.EINVAL
    assert "null handle" in _lib.last_error()


def test_read_index_missing_file_raises_runtimeerror(tmp_path):
    """faiss.read_index on a missing path raises RuntimeError [faiss-upstream]; rag.py:179 turns it
    into HTTP 500.  The file check happens before any device work, so this runs without a GPU."""
    import b200knn
    with pytest.raises(RuntimeError, match="could not open"):
        b200knn.read_index(str(tmp_path / "nope"))
    bad = tmp_path / "bad"
    bad.write_bytes(b"IxFI" + b"\0" * 41)
    with pytest.raises(RuntimeError, match="not supported"):
        b200knn.read_index(str(bad))
    short = tmp_path / "short"
    short.write_bytes(b"IxF2\0\0")
    with pytest.raises(RuntimeError, match="short read"):
        b200knn.read_index(str(short))


def test_default_device_env(monkeypatch):
    import b200knn
    monkeypatch.delenv("B200KNN_DEVICE", raising=False)
    monkeypatch.delenv("LOCAL_RANK", raising=False)
    assert b200knn.default_device() == 0
    monkeypatch.setenv("LOCAL_RANK", "3")
    assert b200knn.default_device() == 3
    monkeypatch.setenv("B200KNN_DEVICE", "1")
    assert b200knn.default_device() == 1


def test_every_option_and_stat_key_is_documented_in_the_header():
    """b2k_set_option / b2k_get_stat take string keys: each key the library accepts must be described in include/b200knn.h."""
    hdr = open(os.path.join(ROOT, "include", "b200knn.h")).read()
    csrc = os.path.join(ROOT, "rag-llm-k8s_b200", "csrc")
    keys = set()
    for f in ("index.cu", "multi.cu"):
        keys |= set(re.findall(r'!strcmp\(key, "(\w+)"\)', open(os.path.join(csrc, f)).read()))
    assert len(keys) >= 20
    missing = sorted(k for k in keys if '"%s"' % k not in hdr)
    assert not missing, "undocumented option / stat keys: %s" % missing
