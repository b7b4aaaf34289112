"""A few searches on one shard-sized index (for a per-kernel launch list under ncu):  shard_one.py [N] [nq] [k] [d] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 1_250_000
nq = a[1] if len(a) > 1 else 1024
k = a[2] if len(a) > 2 else 100
d = a[3] if len(a) > 3 else 384
reps = a[4] if len(a) > 4 else 3
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0)
q = b200knn.synth(5678, 0, nq, d)
for i in range(reps):
    idx.search(q, k)
torch.cuda.synchronize()
