"""Small searches through every kernel family, each checked against the oracle (a quick smoke of all
instantiations; small enough to run under a memory checker where one is available)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import b200knn
from oracle import oracle as O
from util import assert_knn_matches
x = O.synth(1, 0, 90000, 64); 
for opts, nq, k in (({"path": 2, "tensor_mode": 3, "tensor_cta_group": 1}, 70, 10), ({"path": 2, "tensor_mode": 3, "tensor_cta_group": 2}, 300, 10),
                    ({"path": 2, "tensor_mode": 2, "tensor_cta_group": 1, "tensor_epilogue_warps": 4}, 40, 10), ({"path": 2, "tensor_mode": 4}, 40, 10),
                    ({"path": 2, "tensor_mode": 1}, 40, 10), ({"path": 1, "scan_rows": 2}, 3, 10), ({"path": 1, "scan_rows": 1}, 9, 10)):
    idx = b200knn.IndexFlatL2(64); idx.add(x)
    for kk, v in opts.items(): idx.set_option(kk, v)
    q = O.synth(2, 0, nq, 64)
    D, I = idx.search(q, k)
    assert_knn_matches(D, I, x, q, k, expanded=True, what=str(opts))
    print("ok", opts, idx.stats().get("tensor_mode") or idx.stats().get("scan_rows"), flush=True)
# views, the multi-GPU handle (several shards on one GPU through the test hook) and the shard merge
os.environ["B200KNN_TEST_DUP_DEVICES"] = "1"
idx = b200knn.IndexFlatL2(64); idx.add(x[:50000])
v = idx._view(); idx.add(x[50000:]); v.add(x[:10])
q = O.synth(2, 0, 5, 64)
D, I = v.search(q, 10)
assert_knn_matches(D, I, np.concatenate([x[:50000], x[:10]]), q, 10, what="view")
print("ok view", flush=True)
m = b200knn.IndexFlatL2(64, devices=[0, 0, 0]); m.add(x)
for nq in (3, 300):
    q = O.synth(2, 0, nq, 64)
    D, I = m.search(q, 10)
    assert_knn_matches(D, I, x, q, 10, expanded=nq >= 20, what="multi nq=%d" % nq)
print("ok multi", flush=True)
