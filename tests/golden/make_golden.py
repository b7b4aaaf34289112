"""Regenerates the fixtures in this directory:  python tests/golden/make_golden.py

No output of the reference can be produced here (faiss-cpu is absent), so the
expected values come from the float64 brute force in oracle/oracle.py, on inputs drawn from
the counter-based synthetic stream (seeds 1234 / 5678 as in SURVEY.md 8d)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as O  # noqa: E402
from util import unit_rows  # noqa: E402

CASES = {
    "c1_2048x384_nq1_k5": dict(n=2048, d=384, nq=1, k=5, unit=False),
    "app_64x1024_unit_nq1_k5": dict(n=64, d=1024, nq=1, k=5, unit=True),
    "batch_3000x96_nq32_k100": dict(n=3000, d=96, nq=32, k=100, unit=False),
}

for name, c in CASES.items():
    x = O.synth(1234, 0, c["n"], c["d"])
    q = O.synth(5678, 0, c["nq"], c["d"])
    if c["unit"]:
        x, q = unit_rows(x), unit_rows(q)
    Dt, It = O.truth_f64(x, q, c["k"])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), seed_x=1234, seed_q=5678, n=c["n"], d=c["d"],
                        nq=c["nq"], k=c["k"], unit=c["unit"], D=Dt.astype(np.float32), I=It)
    print(name, Dt[0, :3], It[0, :3])

rows = O.synth(42, 7, 4, 8)
np.savez(os.path.join(HERE, "synth_kat.npz"), seed=42, row0=7, rows=rows)
print(rows)
