#!/usr/bin/env python
"""bench.py -- headline benchmark of the IndexFlatL2 search path (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c1|c5|smoke] [--impl reference]

Workload (default c3 = BASELINE.json configs[2], the configuration the metric is quoted on):
10M x 384 fp32 database, batches of nq=1024 queries, k=100.  One "step" = one search batch.
  value  queries/s with database, queries and results resident in HBM (CUDA events);
  e2e    the same through the public API with HOST numpy buffers (H2D of the queries and D2H of
         (D, I) inside the timed region);
  roofline / cpu_baseline as described in DESIGN.md section "Measurement".
With N > 1 (torchrun, one rank per GPU) the 10M rows are row-sharded over the ranks (strong
scaling: total work fixed), each step ends with one NCCL all-gather of the local top-k and the
merge kernel; time is the max over ranks.

`--impl reference` times the CPU restatement of faiss's IndexFlatL2.search (oracle/) on the
host cores on a bounded sample of the same workload.  faiss-cpu itself is not installable here.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

# two OpenMP runtimes meet in the CPU baseline (MKL's and the oracle's libgomp): let idle threads sleep
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
os.environ.setdefault("KMP_BLOCKTIME", "0")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rag-llm-k8s_b200"))

WORKLOADS = {
    # name: (N, d, nq, k)
    "c1": (2048, 384, 1, 5),
    "c2": (1_000_000, 384, 1, 10),
    "c3": (10_000_000, 384, 1024, 100),
    "c3q1": (10_000_000, 384, 1, 100),
    "c4": (100_000_000, 768, 256, 100),      # needs 8 GPUs (38.4 GB of rows + 38.4 GB residual rows per GPU)
    "c5": (10_000_000, 1024, 4096, 1024),
    "smoke": (200_000, 384, 64, 10),
}
SEED_DB, SEED_Q = 1234, 5678


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm=j["hbm_gbs"], bf16=j["bf16_tflops"], bf16_sus=j.get("bf16_tflops_sustained", j["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons while the timed region runs (NVML)."""

    def __init__(self, dev):
        super().__init__(daemon=True)
        self.dev, self.stop_flag, self.sm, self.reasons, self.sm_max = dev, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(dev)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def cpu_reference(N, d, nq, k, steps, warmup, budget_s=15.0):
    """Times the oracle (faiss-cpu-equivalent restatement) on a bounded sample: all nq queries
    against the first S rows; QPS is scaled linearly in N (the scan is linear in rows)."""
    import numpy as np
    from oracle import oracle as O
    threads = O.num_threads()
    if nq >= 20:
        # what faiss does at this size: sgemm (here MKL through torch.mm) + norms/clamp/result handler
        def run(xx):
            return O.search_blas(xx, q, k, db_bs=16384)
        how = "oracle.search_blas: torch.mm (MKL sgemm) blocks + oracle/flat_l2_oracle.c selection, %d threads" % threads
    else:
        def run(xx):
            return O.search(xx, q, k, O.MODE_DIRECT_ROWPAR)
        how = ("oracle/flat_l2_oracle.c direct scan, rows split over %d OpenMP threads "
               "(faiss itself would use 1 thread at nq=1)" % threads)
    q = O.synth(SEED_Q, 0, nq, d)
    S = min(N, 32768)
    x = O.synth(SEED_DB, 0, S, d)
    run(x)
    t0 = time.perf_counter()
    run(x)
    t_cal = max(time.perf_counter() - t0, 1e-4)
    per_step = budget_s / max(steps + warmup, 1)
    S = int(min(N, max(S, S * per_step / t_cal)))
    S = max(min(S, int(8e9 / (4 * d))), 1)
    x = O.synth(SEED_DB, 0, S, d)
    for _ in range(min(warmup, 1)):
        run(x)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        run(x)
        times.append(time.perf_counter() - t0)
    t = sum(times) / len(times) * (N / S)
    return {"value": nq / t, "unit": "queries/s", "cores": threads, "kind": "port",
            "sample": "all %d queries x first %d of %d rows, %d timed passes, time scaled x%.1f (linear in rows); %s" % (
                nq, S, N, steps, N / S, how)}, t * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rows", type=int, default=0, help="override N (debug)")
    ap.add_argument("--tensor-mode", type=int, default=0, choices=[0, 1, 2, 3],
                    help="batched-search numerics: 0 library default, 1 3xTF32, 2 TF32 screening, 3 bf16 screening (all exact results)")
    args = ap.parse_args()

    N, d, nq, k = WORKLOADS[args.workload]
    if args.rows:
        N = args.rows
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    metric = "QPS at %s k=%d nq=%d" % (("%dMx%d" % (N // 1_000_000, d)) if N % 1_000_000 == 0 else "%dx%d" % (N, d), k, nq)
    config = {"workload": "%s: IndexFlatL2 %dx%d fp32, nq=%d, k=%d" % (args.workload, N, d, nq, k),
              "N": N, "d": d, "nq": nq, "k": k, "parallelism": "row-shard x%d" % world if world > 1 else "1 gpu",
              "l2": "database (%.1f GB) is larger than L2, no flush needed" % (N * d * 4 / 1e9),
              "db_seed": SEED_DB, "query_seed": SEED_Q}

    if args.impl == "reference":
        if rank != 0:
            return
        cb, ms = cpu_reference(N, d, nq, k, args.steps, args.warmup, budget_s=60.0)
        line = {"impl": "reference", "metric": metric, "value": cb["value"], "unit": "queries/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "faiss-cpu-equivalent restatement (not faiss-cpu): the wheel cannot be installed here"}
        print(json.dumps(line))
        return

    import numpy as np
    import torch
    import b200knn
    from b200knn import _lib

    if N * d * 8 / max(world, 1) > 150e9:
        raise SystemExit("bench.py: workload %s needs more GPUs (%.0f GB per GPU at --gpus %d)" % (args.workload, N * d * 8 / world / 1e9, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; b200knn has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # ---- build the (shard of the) index on the device -----------------------------------
    if world > 1:
        from b200knn.sharded import ShardedIndexFlatL2
        index = ShardedIndexFlatL2(d, device=local_rank)
        index.add_synthetic_global(N, SEED_DB)
    else:
        index = b200knn.IndexFlatL2(d, device=local_rank)
        index.reserve(N)
        index.add_synthetic(N, SEED_DB, 0)
    index.set_option("timing", 1)
    if args.tensor_mode:
        index.set_option("tensor_mode", args.tensor_mode)
    q_dev = b200knn.synth(SEED_Q, 0, nq, d, device=local_rank)
    # e2e inputs / outputs live in page-locked host memory (the contract's "pinned host memory"): the library
    # then copies to / from them directly instead of staging through its own pinned buffer
    q_pin = torch.empty(q_dev.shape, dtype=torch.float32, pin_memory=True)
    q_pin.copy_(q_dev)
    q_host = q_pin.numpy()
    D_pin = torch.empty((nq, k), dtype=torch.float32, pin_memory=True).numpy()
    I_pin = torch.empty((nq, k), dtype=torch.int64, pin_memory=True).numpy()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = b200knn.launch_count()
        t0 = time.perf_counter()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms = ev0.elapsed_time(ev1)
        launches = b200knn.launch_count() - l0
        if dist is not None:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall * 1e3, launches

    # device-resident arm (value) and host-buffer arm (e2e)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev, _, launches = timed(lambda: index.search(q_dev, k), args.steps, args.warmup)
    sampler.stop_flag = True
    sampler.join()
    stats = index.stats() if hasattr(index, "stats") else {}
    if world > 1:
        e2e_call = lambda: index.search(q_host, k)            # (the sharded index allocates its own outputs)
    else:
        e2e_call = lambda: index.search(q_host, k, D=D_pin, I=I_pin)
    _, wall_e2e, _ = timed(e2e_call, args.steps, max(args.warmup, 1))
    if dist is not None:
        t = torch.tensor([wall_e2e], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall_e2e = float(t.item())

    # p50 single-query latency on the same index (second half of BASELINE.json's metric):
    # CUDA events around the public call (includes the host's submission gap) and the scan
    # kernel alone (events recorded inside the library around the launch)
    q1 = q_dev[:1].contiguous()
    lat, lat_k = [], []
    nq1_rows = "fp32"
    for i in range(30):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        index.search(q1, k)
        e1.record()
        torch.cuda.synchronize()
        if i >= 5:
            lat.append(e0.elapsed_time(e1))
            st1 = index.stats()
            if st1.get("path") == "scan" and st1.get("main_kernel_ms"):
                lat_k.append(st1["main_kernel_ms"])
                nq1_rows = st1.get("scan_rows", "fp32")
    p50_nq1 = statistics.median(lat)
    p50_nq1_kernel = statistics.median(lat_k) if lat_k else None
    ms_step = ms_dev / args.steps
    qps = nq / (ms_step * 1e-3)
    qps_e2e = nq / (wall_e2e / args.steps * 1e-3)

    # ---- roofline of the dominant kernel ----------------------------------------------
    pk = peaks()
    tf32_meas = None
    if nq >= 20 and rank == 0:
        # cuBLAS TF32 GEMM on this box (MEASURED_PEAKS.json has no TF32 row): best of 5 and a 1.5 s loop
        torch.backends.cuda.matmul.allow_tf32 = True
        a = torch.randn(8192, 8192, device="cuda"); b = torch.randn(8192, 8192, device="cuda")
        for _ in range(2):
            a @ b
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(int(1500.0 / best), 1)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record(); torch.cuda.synchronize()
        tf32_meas = {"burst": 2 * 8192 ** 3 / (best * 1e-3) / 1e12,
                     "sustained": 2 * 8192 ** 3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12,
                     "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), CUDA events"}
        del a, b
        torch.backends.cuda.matmul.allow_tf32 = False
    n_local = N // world if world > 1 else N
    if stats.get("path") == "tensor":
        k_ms = stats["main_kernel_ms"]
        tmode = stats.get("tensor_mode")
        bf16 = tmode == "bf16screen+rerank"
        screen = bf16 or tmode == "screen+rerank"
        nprod = 1 if screen else 3
        flops = nprod * 2.0 * nq * n_local * d
        peak = pk["bf16_sus"] if bf16 else pk["bf16_sus"] / 2.0
        roof = {"bound": "tensor", "kernel": "k3_tf32x3_topk", "tensor_mode": tmode,
                "screen_fallbacks": stats.get("screen_fallbacks"),
                "achieved": flops / (k_ms * 1e-3) / 1e12, "peak": peak,
                "unit": "TFLOP/s", "traffic": None, "launch_ms": k_ms,
                "peak_source": ("%s bf16 sustained (dense bf16 MMA)" % pk["source"]) if bf16 else
                "%s bf16 sustained / 2 (TF32 dense; derived, MEASURED_PEAKS.json has no TF32 row)" % pk["source"],
                "algorithmic": "executed tensor flops = %d * 2*nq*N*d per launch (%s)" % (
                    nprod, ("one bf16 product" if bf16 else "one TF32 product") +
                    "; exact distances come from the fp32 re-rank of the rows inside its error slack"
                    if screen else "3xTF32"),
                "fp32_equivalent_tflops": 2.0 * nq * n_local * d / (k_ms * 1e-3) / 1e12, "tf32_cublas_measured": tf32_meas}
    else:
        per_launch_ms = (stats["main_kernel_ms"] / max(stats.get("launches", 1), 1)) if stats.get("main_kernel_ms") \
            else ms_dev / max(launches, 1)
        bpe = 2 if stats.get("scan_rows") == "bf16+rerank" else 4
        gbs = n_local * d * bpe / (per_launch_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "scan_l2_topk_kernel", "scan_rows": stats.get("scan_rows", "fp32"),
                "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                "traffic": None, "launch_ms": per_launch_ms, "peak_source": "%s copy bandwidth" % pk["source"],
                "algorithmic": "N*d*%d bytes per launch (one pass over the %s per group of queries%s)" % (
                    bpe, "bf16 copies of the rows" if bpe == 2 else "fp32 rows",
                    "; the k best come from an exact fp32 re-rank of 2k+32 candidates" if bpe == 2 else "")}
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        tj = json.load(open(tr)).get("%s:%s:n%d" % (args.workload, roof["kernel"], world))
        if tj:
            roof["traffic"] = tj["bytes_per_launch"]
            roof["traffic_source"] = tj["source"]
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["scope"] = "one GPU (rank 0's shard of %d rows)" % n_local if world > 1 else "the one GPU"
    # single-query scan (the HBM-bound path) on the same index
    t_nq1 = p50_nq1_kernel or p50_nq1
    bpe1 = 2 if nq1_rows == "bf16+rerank" else 4
    roof_nq1 = {"bound": "hbm", "kernel": "scan_l2_topk_kernel", "scan_rows": nq1_rows,
                "achieved": n_local * d * bpe1 / (t_nq1 * 1e-3) / 1e9,
                "algorithmic": "N*d*%d bytes (one pass over the %s)" % (bpe1, "bf16 copies of the rows" if bpe1 == 2 else "fp32 rows"),
                "peak": pk["hbm"], "unit": "GB/s", "launch_ms": t_nq1,
                "timed": "kernel (in-library CUDA events)" if p50_nq1_kernel else "public call (CUDA events)"}
    roof_nq1["frac"] = roof_nq1["achieved"] / roof_nq1["peak"]

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    cb = None
    if world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_reference(N, d, nq, k, 2, 1, budget_s=15.0)
    line = {"metric": metric, "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "e2e": {"value": qps_e2e, "unit": "queries/s", "h2d_bytes_per_step": nq * d * 4,
                    "d2h_bytes_per_step": nq * k * 12},
            "gpu_launches": launches, "clocks": sampler.result(), "roofline": roof,
            "p50_nq1_ms": p50_nq1, "roofline_nq1": roof_nq1, "cpu_baseline": cb}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
