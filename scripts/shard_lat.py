"""Per-phase timing of the sharded nq=1 search (run under torchrun)."""
import os, sys, statistics, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, torch.distributed as dist
import b200knn
from b200knn.sharded import ShardedIndexFlatL2
lr = int(os.environ.get("LOCAL_RANK", "0")); torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N, d, k = 10_000_000, 384, 100
idx = ShardedIndexFlatL2(d, device=lr); idx.add_synthetic_global(N, 1234); idx.set_option("timing", 1)
q = b200knn.synth(5678, 0, 1, d, device=lr)
for nq in (1, 1024):
    qq = b200knn.synth(5678, 0, nq, d, device=lr)
    ev, wall, kern = [], [], []
    for i in range(25):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(); idx.search(qq, k); e1.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
        if i >= 5:
            ev.append(e0.elapsed_time(e1)); wall.append((t1 - t0) * 1e3); kern.append(idx.stats()["main_kernel_ms"])
    if dist.get_rank() == 0:
        print("nq=%d world=%d: events %.3f ms, wall %.3f ms, local kernel %.3f ms" % (nq, dist.get_world_size(), statistics.median(ev), statistics.median(wall), statistics.median(kern)), flush=True)
dist.destroy_process_group()
