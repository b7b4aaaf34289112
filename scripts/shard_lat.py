"""Per-phase timing of the sharded search (run under torchrun):  shard_lat.py [N] [nq] [k] [d]"""
import os, sys, statistics, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))
import torch, torch.distributed as dist
import b200knn
from b200knn.sharded import ShardedIndexFlatL2
lr = int(os.environ.get("LOCAL_RANK", "0")); torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
a = [int(v) for v in sys.argv[1:]]
N = a[0] if len(a) > 0 else 10_000_000
nq = a[1] if len(a) > 1 else 1024
k = a[2] if len(a) > 2 else 100
d = a[3] if len(a) > 3 else 384
idx = ShardedIndexFlatL2(d, device=lr); idx.add_synthetic_global(N, 1234); idx.set_option("timing", 1)
qq = b200knn.synth(5678, 0, nq, d, device=lr)
W = dist.get_world_size()


def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e


rows = []
from b200knn import merge_shards
for i in range(25):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    nk, words, (send, recv, slot, p2p) = idx._buffers(nq, k)
    I_loc = send[:nk].view(nq, k); D_loc = send[nk:words].view(torch.float32)[:nk].view(nq, k)
    e0 = ev()
    send[words:].zero_()
    idx.local.search(qq, k, D=D_loc, I=I_loc)
    e1 = ev()
    e2 = ev()
    rnccl = torch.empty(W * slot, dtype=torch.int64, device=qq.device) if i == 0 else rnccl
    dist.all_gather_into_tensor(rnccl, send)          # (the NCCL form of the exchange, timed for comparison)
    e3 = ev()
    Dm = torch.empty((nq, k), dtype=torch.float32, device=qq.device); Im = torch.empty((nq, k), dtype=torch.int64, device=qq.device)
    merge_shards(rnccl.data_ptr() + 8 * nk, 2 * slot, rnccl.data_ptr(), slot, W, nq, k, Dm, Im, cyclic_block=idx.block)
    e4 = ev()
    torch.cuda.synchronize(); t1 = time.perf_counter()
    if i >= 5:
        rows.append((e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3), e3.elapsed_time(e4), (t1 - t0) * 1e3,
                     idx.local.stats().get("main_kernel_ms", 0.0)))
    # whole call through the public method, for comparison
tot = []
for i in range(15):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0 = ev(); idx.search(qq, k); e1 = ev(); torch.cuda.synchronize()
    if i >= 5: tot.append(e0.elapsed_time(e1))
if dist.get_rank() == 0:
    med = [statistics.median(c) for c in zip(*rows)]
    print(("N=%d nq=%d k=%d d=%d world=%d exchange=" + ("p2p" if idx._p2p is not None else "nccl") + ": local search %.3f (main pass %.3f) | pack %.3f | NCCL all_gather %.3f | merge %.3f | wall %.3f ms ; public search() %.3f ms")
          % (N, nq, k, d, W, med[0], med[5], med[1], med[2], med[3], med[4], statistics.median(tot)), flush=True)
dist.destroy_process_group()
