"""Where the batched pass spends its time: probe_levels.py [N] [nq] [k] [d] -- main-pass time with the epilogue
progressively switched off (option tensor_probe: 0 full, 1 filter shut, 2 accumulator loads only, 3 MMA pipeline alone),
for the streaming and the query-stationary schedule."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 10_000_000
nq = a[1] if len(a) > 1 else 1024
k = a[2] if len(a) > 2 else 100
d = a[3] if len(a) > 3 else 384
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
for mode in (3, 2):
    idx.set_option("tensor_mode", mode)
    for qs in (0, 1):
        idx.set_option("tensor_query_stationary", qs)
        for probe in (0, 1, 2, 3):
            idx.set_option("tensor_probe", probe)
            main = []
            for i in range(8):
                idx.search(q, k); torch.cuda.synchronize()
                if i >= 3: main.append(idx.stats()["main_kernel_ms"])
            ms = statistics.median(main)
            print("N=%d nq=%d d=%d mode=%d qs=%d probe=%d: main pass %.3f ms = %.0f TFLOP/s" % (N, nq, d, mode, qs, probe, ms, 2.0 * nq * N * d / (ms * 1e-3) / 1e12), flush=True)
idx.set_option("tensor_probe", 0)
