"""IxF2 file format known answers (CPU).  [faiss-upstream]: layout recorded in SURVEY.md 8a-a5,
unverified against a real faiss build."""
import numpy as np
import pytest

from oracle import oracle as O


@pytest.mark.parametrize("d", [1, 384, 1024])
def test_empty_index_is_the_45_known_bytes(tmp_path, d):
    p = tmp_path / "idx"
    O.write_ixf2(str(p), np.zeros((0, d), np.float32), d=d)
    raw = p.read_bytes()
    assert len(raw) == 45 and raw == O.empty_ixf2_bytes(d)
    dd, x = O.read_ixf2(str(p))
    assert dd == d and x.shape == (0, d)


def test_size_and_round_trip(tmp_path):
    x = O.synth(9, 0, 37, 12)
    p = tmp_path / "idx"
    O.write_ixf2(str(p), x)
    raw = p.read_bytes()
    assert len(raw) == 45 + 4 * 37 * 12
    assert raw[:4] == b"IxF2" and int.from_bytes(raw[8:16], "little") == 37
    assert int.from_bytes(raw[37:45], "little") == 37 * 12
    d, y = O.read_ixf2(str(p))
    assert d == 12 and np.array_equal(x, y)
