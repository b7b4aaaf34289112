"""Sensitivity of the batched pass to the depth of its operand ring: k <= 256 runs 6 stages, 257..768 run 5, longer lists 4 (room for
longer lists); with the epilogue off (tensor_probe 3) the difference is the pipeline's alone."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
N, nq, d = 10_000_000, 1024, 384
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
for rep in range(2):
    for k in (100, 300):
        for probe in (3, 0):
            idx.set_option("tensor_probe", probe)
            main = []
            for i in range(8):
                idx.search(q, k); torch.cuda.synchronize()
                if i >= 3: main.append(idx.stats()["main_kernel_ms"])
            ms = statistics.median(main)
            print("k=%d (%d stages) probe=%d: main pass %.3f ms = %.0f TFLOP/s" % (k, 6 if k <= 256 else 5 if k <= 768 else 4, probe, ms, 2.0 * nq * N * d / (ms * 1e-3) / 1e12), flush=True)
