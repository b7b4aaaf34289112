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

Before anything is timed the batch's answer is checked (`parity_check` in the JSON line; a mismatch
exits non-zero): a sample of the queries, spread over every query block, against the CPU oracle
streaming over the full row set -- or, where the rows cannot be regenerated on the host in
reasonable time (c4), against the exact scan kernel run per shard and merged by the oracle.
After the headline timing, outside it: the 3xTF32 contraction the north_star names
(`roofline_3xtf32`, against a TF32 peak measured on this box), the single-query scan over the fp32
rows (`roofline_nq1_fp32`), and nq = 1 through host numpy as llm/rag.py:116 calls it
(`p50_nq1_host_ms`).  At --gpus 8 the run ends with BASELINE.json configs[3] (100M x 768 over the
8 shards, nq = 256) in a `c4` block.

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
    "c4": (100_000_000, 768, 256, 100),      # needs 8 GPUs: 38.4 GB of rows + 19.2 GB of bf16 copies per GPU
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
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
                 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        self.power, self.mask = [], 0
        try:
            self.power_limit = nv.nvmlDeviceGetEnforcedPowerLimit(self.h) / 1000.0
        except Exception:
            self.power_limit = None
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.mask |= int(r)
                for bit, name in names.items():
                    if (r & bit) and name != "gpu_idle":
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.002)

    def result(self):
        out = {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
               "reasons": sorted(self.reasons), "samples": len(self.sm)}
        if self.nv:
            out["reasons_mask"] = "0x%x" % getattr(self, "mask", 0)
            pw = getattr(self, "power", [])
            out["power_w"] = statistics.median(pw) if pw else None
            out["power_limit_w"] = getattr(self, "power_limit", None)
            if self.sm and not self.reasons and out["sm_mhz"] < 0.9 * (self.sm_max or 0):
                out["note"] = ("SM clock below its maximum under load with no throttle reason flagged: the board's power management "
                               "under sustained tensor-core load (MEASURED_PEAKS.json records the same for cuBLAS: "
                               "clocks_under_load.sm_mhz_median well below sm_max_mhz, throttled false); no clock lock was set")
        return out


def cpu_reference(N, d, nq, k, steps, warmup, budget_s=15.0):
    """Times the oracle (faiss-cpu-equivalent restatement) on a bounded sample: all nq queries
    against the first S rows; QPS is scaled linearly in N (the scan is linear in rows)."""
    import numpy as np
    from oracle import oracle as O
    threads = O.use_all_cores()      # torchrun exports OMP_NUM_THREADS=1 to its workers
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
            "extrapolated": S < N, "rows_timed": S,
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
    ap.add_argument("--c4-rows", type=int, default=0, help="also run the c4 block (d=768, nq=256) with this many rows at any --gpus (debug)")
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

    if N * d * 6 / max(world, 1) > 165e9:
        raise SystemExit("bench.py: workload %s needs more GPUs (%.0f GB per GPU at --gpus %d)" % (args.workload, N * d * 6 / world / 1e9, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; b200knn has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = dict(np=np, torch=torch, b200knn=b200knn, dist=dist, rank=rank, world=world, local_rank=local_rank, args=args)

    res = run_workload(ctx, args.workload, N, d, nq, k, args.steps, args.warmup, extras=True, config=config)
    c4 = None
    if (world == 8 and args.workload == "c3" and not args.rows) or args.c4_rows:
        # BASELINE.json configs[3] needs all eight GPUs: measured here, where the driver's scaling run looks
        Nc, dc, nqc, kc = WORKLOADS["c4"]
        Nc = args.c4_rows or Nc
        r4 = run_workload(ctx, "c4", Nc, dc, nqc, kc, args.steps, args.warmup, extras=False)
        if rank == 0:
            c4 = {"workload": "c4: IndexFlatL2 %dx%d fp32 row-sharded x%d, nq=%d, k=%d" % (Nc, dc, world, nqc, kc),
                  "qps": r4["qps"], "ms_per_step": r4["ms_step"], "e2e": r4["e2e"], "roofline": r4["roofline"],
                  "parity_check": r4["parity_check"], "p50_nq1_ms": r4["p50_nq1_ms"], "roofline_nq1": r4["roofline_nq1"],
                  "clocks": r4["clocks"], "gpu_launches": r4["launches"],
                  "note": "per-GPU shard %d x 768; shard_main_pass_qps = nq / roofline.launch_ms is what one GPU's main pass "
                          "alone would allow for the whole batch" % (Nc // world),
                  "shard_main_pass_qps": nqc / (r4["roofline"]["launch_ms"] * 1e-3) if r4["roofline"].get("launch_ms") else None}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    cb = None
    if world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_reference(N, d, nq, k, 2, 1, budget_s=15.0)
    line = {"metric": metric, "value": res["qps"], "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "e2e": res["e2e"], "gpu_launches": res["launches"], "clocks": res["clocks"], "roofline": res["roofline"],
            "parity_check": res["parity_check"],
            "p50_nq1_ms": res["p50_nq1_ms"], "p50_nq1_host_ms": res.get("p50_nq1_host_ms"), "roofline_nq1": res["roofline_nq1"],
            "roofline_nq1_fp32": res.get("roofline_nq1_fp32"), "roofline_3xtf32": res.get("roofline_3xtf32"),
            "tf32_peak": res.get("tf32_peak"), "cpu_baseline": cb}
    if c4 is not None:
        line["c4"] = c4
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def parity_check(ctx, index, q_dev, N, d, nq, k, nsel=16):
    """Checks the batch's answer before anything is timed.  Every rank searches; rank 0 compares a
    sample of the queries with the oracle (the checker -- never the thing measured)."""
    np, torch, dist, rank, world = ctx["np"], ctx["torch"], ctx["dist"], ctx["rank"], ctx["world"]
    from oracle import oracle as O
    D, I = index.search(q_dev, k)
    torch.cuda.synchronize()
    nsel = min(nsel, nq)
    sel = np.unique(np.linspace(0, nq - 1, nsel).round().astype(np.int64))     # spread over every query block
    sel_t = torch.as_tensor(sel, device=q_dev.device)
    Dg, Ig = D[sel_t].cpu().numpy(), I[sel_t].cpu().numpy()
    qh = q_dev[sel_t].cpu().numpy()
    expanded = nq >= 20   # faiss's rule for the whole batch (rag.py:116): norm expansion from 20 queries on
    streaming = N * d <= 12e9
    out = None
    if streaming:
        if rank == 0:
            O.use_all_cores()
            t0 = time.perf_counter()
            Dref, Iref = O.stream_search(lambda j0, nb: O.synth(SEED_DB, j0, nb, d), N, d, qh, k,
                                         mode=O.MODE_EXPAND if expanded else O.MODE_DIRECT)
            out = O.compare_knn(Dg, Ig, Dref, Iref, O.synth_rows(SEED_DB, d), qh, rtol=1e-5, expanded=expanded)
            out.update(queries=int(len(sel)), against="CPU oracle streaming over all %d rows (%s rule), %.1f s" % (
                N, "norm-expansion" if expanded else "direct-difference", time.perf_counter() - t0))
    else:
        # the rows cannot be regenerated on the host in reasonable time: every shard answers the sample with the
        # exact scan kernel (direct fp32 differences, path=1), the oracle merges the shard lists
        loc = index.local if world > 1 else index
        loc.set_option("path", 1)
        loc.set_option("scan_rows", 1)
        Ds, Is = loc.search(q_dev[sel_t].contiguous(), k)
        loc.set_option("path", 0)
        loc.set_option("scan_rows", 0)
        if world > 1:
            Is = index.to_global(Is)
            Dl = [torch.empty_like(Ds) for _ in range(world)]
            Il = [torch.empty_like(Is) for _ in range(world)]
            dist.all_gather(Dl, Ds.contiguous())
            dist.all_gather(Il, Is.contiguous())
            Ds, Is = torch.stack(Dl), torch.stack(Il)
        else:
            Ds, Is = Ds[None], Is[None]
        if rank == 0:
            Dref, Iref = O.merge(Ds.cpu().numpy(), Is.cpu().numpy())
            out = O.compare_knn(Dg, Ig, Dref, Iref, O.synth_rows(SEED_DB, d), qh, rtol=1e-5, expanded=False)
            out.update(queries=int(len(sel)), against="exact scan kernel (path=1, fp32 rows) per shard + oracle merge of the "
                                                      "%d shard lists; float64 distances of differing ids from regenerated rows" % world)
    if dist is not None:
        flag = torch.tensor([0 if (out is None or out["ok"]) else 1], device="cuda")
        dist.all_reduce(flag)
        bad = bool(flag.item())
    else:
        bad = not out["ok"]
    if bad:
        if rank == 0:
            print(json.dumps({"parity_check": out, "error": "results differ from the oracle; nothing was timed"}))
        if dist is not None:
            dist.destroy_process_group()
        raise SystemExit(3)
    return out


def tf32_peaks(ctx, index, q_dev, k):
    """TF32 dense rates measured on this box: cuBLAS (torch.matmul with allow_tf32) and this library's own
    kind::tf32 pass with the selection filter shut (option tensor_probe)."""
    torch = ctx["torch"]
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
    out = {"cublas_burst": 2 * 8192 ** 3 / (best * 1e-3) / 1e12,
           "cublas_sustained": 2 * 8192 ** 3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12,
           "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), CUDA events, best of 5 / 1.5 s loop; "
                  "own_dense_pass(_3products) = this library's kind::tf32 pass over the bench index with one (three) product(s) per "
                  "k-step and the selection filter shut (executed flop / in-library CUDA-event time), best of 3"}
    del a, b
    torch.backends.cuda.matmul.allow_tf32 = False
    return out


def run_workload(ctx, name, N, d, nq, k, steps, warmup, extras, config=None):
    config = config if config is not None else {}
    np, torch, b200knn, dist = ctx["np"], ctx["torch"], ctx["b200knn"], ctx["dist"]
    rank, world, local_rank = ctx["rank"], ctx["world"], ctx["local_rank"]
    args = ctx["args"]

    # ---- build the (shard of the) index on the device -----------------------------------
    if world > 1:
        from b200knn.sharded import ShardedIndexFlatL2
        index = ShardedIndexFlatL2(d, device=local_rank)
        index.add_synthetic_global(N, SEED_DB)
        n_local = index.local.ntotal
    else:
        index = b200knn.IndexFlatL2(d, device=local_rank)
        index.reserve(N)
        index.add_synthetic(N, SEED_DB, 0)
        n_local = N
    index.set_option("timing", 1)
    if world > 1:
        config["exchange"] = "one-shot all-gather over peer memory (b2k_exchange_push)" if index._p2p is not None else "NCCL all_gather_into_tensor"
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
            t = torch.tensor([ms, wall * 1e3], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0].item()), float(t[1].item()) * 1e-3
        return ms, wall * 1e3, launches

    pc = parity_check(ctx, index, q_dev, N, d, nq, k)

    # device-resident arm (value) and host-buffer arm (e2e)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev, _, launches = timed(lambda: index.search(q_dev, k), steps, warmup)
    stats = index.stats() if hasattr(index, "stats") else {}
    e2e_call = lambda: index.search(q_host, k, D=D_pin, I=I_pin)
    _, wall_e2e, _ = timed(e2e_call, steps, max(warmup, 1))
    sampler_extended = False
    few = 1 if len(sampler.sm) < 5 else 0
    if dist is not None:             # the sharded search is collective: every rank must take the same decision
        t = torch.tensor([few], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        few = int(t.item())
    if few:
        # the timed regions are tens of milliseconds and an NVML query can take as long when several ranks ask at
        # once: keep the same load running (untimed, ~0.3 s, the same count on every rank) so that the sampler
        # gets a handful of readings under load
        sampler_extended = True
        for _ in range(int(min(400, max(10, 300.0 / (ms_dev / steps))))):
            index.search(q_dev, k)
        torch.cuda.synchronize()
    sampler.stop_flag = True          # (sampled through both timed regions: they are tens of milliseconds long)
    sampler.join()

    # p50 single-query latency on the same index (second half of BASELINE.json's metric):
    # CUDA events around the public call (includes the host's submission gap) and the scan
    # kernel alone (events recorded inside the library around the launch)
    q1 = q_dev[:1].contiguous()

    def nq1_latency(iters=30):
        lat, lat_k, rows = [], [], "fp32"
        for i in range(iters):
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
                    rows = st1.get("scan_rows", "fp32")
        return statistics.median(lat), (statistics.median(lat_k) if lat_k else None), rows

    def nq1_roofline(p50, p50_kernel, rows):
        t = p50_kernel or p50
        bpe = 2 if rows == "bf16+rerank" else 4
        r = {"bound": "hbm", "kernel": "scan_l2_topk_kernel", "scan_rows": rows,
             "achieved": n_local * d * bpe / (t * 1e-3) / 1e9,
             "algorithmic": "N*d*%d bytes (one pass over the %s)" % (bpe, "bf16 copies of the rows" if bpe == 2 else "fp32 rows"),
             "peak": pk["hbm"], "unit": "GB/s", "launch_ms": t, "p50_call_ms": p50,
             "timed": "kernel (in-library CUDA events)" if p50_kernel else "public call (CUDA events)"}
        r["frac"] = r["achieved"] / r["peak"]
        return r

    pk = peaks()
    p50_nq1, p50_nq1_kernel, nq1_rows = nq1_latency()
    roof_nq1 = nq1_roofline(p50_nq1, p50_nq1_kernel, nq1_rows)
    ms_step = ms_dev / steps
    qps = nq / (ms_step * 1e-3)
    qps_e2e = nq / (wall_e2e / steps * 1e-3)

    # ---- roofline of the dominant kernel ----------------------------------------------
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
                "%s bf16 sustained / 2 (see tf32_peak for the rates measured in this run)" % pk["source"],
                "frac_of_burst_peak": (flops / (k_ms * 1e-3) / 1e12) / pk["bf16"] if bf16 else None,
                "algorithmic": "executed tensor flops = %d * 2*nq*N*d per launch (%s)" % (
                    nprod, ("one bf16 product" if bf16 else "one TF32 product") +
                    "; exact distances come from the fp32 re-rank of the rows inside its error slack"
                    if screen else "3xTF32"),
                "fp32_equivalent_tflops": 2.0 * nq * n_local * d / (k_ms * 1e-3) / 1e12}
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
        tj = json.load(open(tr)).get("%s:%s:n%d" % (name, roof["kernel"], 1))
        if tj:
            scale = n_local / float(N)      # a shard streams its share of the rows
            roof["traffic"] = tj["bytes_per_launch"] * scale
            roof["traffic_source"] = tj["source"] + ("" if world == 1 else " (1-GPU capture scaled by the shard's share of the rows)")
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["scope"] = "one GPU (rank 0's shard of %d rows)" % n_local if world > 1 else "the one GPU"

    clocks = sampler.result()
    if sampler_extended:
        clocks["extended"] = "the same searches ran on, untimed, for ~0.3 s so that NVML could be read under load"
    out = {"qps": qps, "ms_step": ms_step, "launches": launches, "clocks": clocks, "roofline": roof,
           "parity_check": pc, "p50_nq1_ms": p50_nq1, "roofline_nq1": roof_nq1,
           "e2e": {"value": qps_e2e, "unit": "queries/s", "h2d_bytes_per_step": nq * d * 4, "d2h_bytes_per_step": nq * k * 12}}

    if extras:
        loc = index.local if world > 1 else index
        # nq = 1 exactly as llm/rag.py:116 makes the call: a pageable numpy query, fresh numpy results, host clock
        qh1 = np.ascontiguousarray(q_host[:1]).copy()
        hl = []
        for i in range(30):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            index.search(qh1, k)
            hl.append((time.perf_counter() - t0) * 1e3)
        out["p50_nq1_host_ms"] = statistics.median(hl[5:])
        # north_star: ">= 80% of the HBM roofline at nq = 1" refers to the fp32 rows (N*d*4 bytes)
        loc.set_option("scan_rows", 1)
        a, b, rows = nq1_latency(20)
        out["roofline_nq1_fp32"] = nq1_roofline(a, b, rows)
        loc.set_option("scan_rows", 0)
        if nq >= 20 and not args.tensor_mode:
            # north_star: ">= 60% of the 3xTF32 tensor-core roofline at nq = 1024": the contraction with three
            # TF32 products per k-step (tensor_mode = 1), against the TF32 dense rate measured on this box
            tp = tf32_peaks(ctx, index, q_dev, k) if rank == 0 else {}
            barrier()          # rank 0 has just spent seconds on its own: the ranks go on together
            for pm, key, nprod in ((2, "own_dense_pass", 1), (1, "own_dense_pass_3products", 3)):
                probe_ms = None
                try:
                    loc.set_option("tensor_mode", pm)
                    loc.set_option("tensor_probe", 1)
                    for i in range(4):
                        index.search(q_dev, k)
                        torch.cuda.synchronize()
                        ms_i = index.stats().get("main_kernel_ms")
                        if i >= 1 and ms_i:
                            probe_ms = ms_i if probe_ms is None else min(probe_ms, ms_i)
                finally:
                    loc.set_option("tensor_probe", 0)
                    loc.set_option("tensor_mode", 0)
                if probe_ms:
                    tp[key] = nprod * 2.0 * nq * n_local * d / (probe_ms * 1e-3) / 1e12
                    tp[key + "_ms"] = probe_ms
            loc.set_option("tensor_mode", 1)
            try:
                ms3, _, _ = timed(lambda: index.search(q_dev, k), 3, 2)
                st3 = index.stats()
            finally:
                loc.set_option("tensor_mode", 0)
            if st3.get("main_kernel_ms") and rank == 0:
                k3ms = st3["main_kernel_ms"]
                executed = 3 * 2.0 * nq * n_local * d / (k3ms * 1e-3) / 1e12
                tp["derived_bf16_burst_over_2"] = pk["bf16"] / 2.0
                cands = {kk: tp[kk] for kk in ("cublas_burst", "cublas_sustained", "own_dense_pass", "own_dense_pass_3products",
                                               "derived_bf16_burst_over_2") if tp.get(kk)}
                peak_name = max(cands, key=cands.get)
                peak3 = cands[peak_name]
                out["roofline_3xtf32"] = {
                    "bound": "tensor", "kernel": "k3_tf32x3_topk", "tensor_mode": st3.get("tensor_mode"),
                    "launch_ms": k3ms, "ms_per_step": ms3 / 3, "qps": nq / (ms3 / 3 * 1e-3),
                    "achieved": executed, "unit": "TFLOP/s", "peak": peak3, "frac": executed / peak3,
                    "peak_source": "tf32_peak.%s: the largest of the TF32 dense rates measured in this run (cuBLAS burst / sustained, this "
                                   "library's kind::tf32 passes with the filter shut) and %s bf16 burst / 2 (derived; it is the ceiling "
                                   "whenever no TF32 run reaches it)" % (peak_name, pk["source"]),
                    "algorithmic": "SURVEY.md 8(d): 3 * 2*nq*N*d executed tensor flops per batch",
                    "fp32_equivalent_tflops": executed / 3}
            out["tf32_peak"] = tp
    del index
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    main()
