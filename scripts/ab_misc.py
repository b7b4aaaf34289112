import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, b200knn
N, nq, d, k = 10_000_000, 1024, 384, 100
idx = b200knn.IndexFlatL2(d); idx.reserve(N); idx.add_synthetic(N, 1234, 0); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, nq, d)
def run(tag, n=8):
    main = []
    for i in range(n):
        idx.search(q, k); torch.cuda.synchronize()
        if i >= 3: main.append(idx.stats()["main_kernel_ms"])
    print("%s: main pass %.3f ms (min %.3f max %.3f)" % (tag, statistics.median(main), min(main), max(main)), flush=True)
for rep in range(3):
    for probe in (3, 2, 1, 0):
        idx.set_option("tensor_probe", probe)
        for tm in (0, 3):
            idx.set_option("tensor_mode", tm)
            run("rep %d probe=%d tensor_mode=%d" % (rep, probe, tm))
