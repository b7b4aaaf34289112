"""Worst-case accuracy of the tensor-core path: unit-norm rows, queries equal to stored rows
(true distance 0, accumulator grows monotonically to 1).  Error reported relative to
||q||^2 + ||x||^2 = 2, for each accumulator mode."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import b200knn
from oracle import oracle as O
from util import unit_rows
for d in (128, 384, 512, 768, 1024, 2048):
    x = unit_rows(O.synth(1, 0, 4096, d)); q = x[:64].copy()
    out = []
    for sep, name in ((2, "one-accum"), (1, "sep-accum")):
        for cg in (1, 2):
            idx = b200knn.IndexFlatL2(d); idx.add(x); idx.set_option("path", 2)
            idx.set_option("tensor_cta_group", cg); idx.set_option("tensor_sep_accum", sep)
            D, I = idx.search(q, 1)
            assert np.array_equal(I[:, 0], np.arange(64)), (d, sep, cg)
            out.append("%s/cg%d %.2e" % (name, cg, float(D.max()) / 2.0))
    print("d=%d  max |D - 0| / (|q|^2+|x|^2):  %s" % (d, "  ".join(out)), flush=True)
