"""Turn the ncu outputs in gpurun_out/ into the tracked summaries under profiles/.

    python scripts/summarise_profiles.py r01
"""
import csv
import json
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 else "r01"
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)


def launches(path):
    rows = list(csv.reader(open(path)))
    for i, r in enumerate(rows):
        if r and r[0] == "ID":
            hdr, start = r, i + 1
            break
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    out = []
    for r in rows[start:]:
        if len(r) > vi:
            v = float(r[vi].replace(",", ""))
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
            out.append((re.sub(r"\(.*", "", r[ki]).replace("void ", "").strip(), v))
    return out


def raw_metrics_all(rep, wanted):
    """One OrderedDict per captured launch."""
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    out = []
    for vals in rows[2:]:
        res = OrderedDict()
        for w in wanted:
            if w in hdr:
                i = hdr.index(w)
                res[w] = (vals[i], units[i])
        out.append(res)
    return out


def raw_metrics(rep, wanted):
    return raw_metrics_all(rep, wanted)[0]


WANTED = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
          "l1tex__m_xbar2l1tex_read_bytes.sum",
          "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
          "sm__cycles_elapsed.avg.per_second", "sm__warps_active.avg.pct_of_peak_sustained_active",
          "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
          "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum"]


def to_bytes(v, u):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]


traffic = {}
md = ["# Round %s profile summary (B200, `scripts/profile.sh %s`; raw files in gpurun_out/, not tracked)\n" % (R, R)]
def to_ms(v, u):
    return float(v.replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[u]


# K3: the main pass of one search is a handful of launches (segments of the row range with a
# bound refresh in between, after one small warm-start launch): all of them are captured; the
# table shows the largest segment, the traffic is summed over the segments.
k3_rep = os.path.join(G, "%s_k3_c3.ncu-rep" % R)
if os.path.exists(k3_rep):
    allm = raw_metrics_all(k3_rep, WANTED)
    # launch 0 = warm start; the segments of ONE main pass read the bf16 rows once (7.68 GB at C3, or the
    # 15.36 GB of fp32 rows in the TF32 modes): the capture count is fixed, the number of segments is
    # not, so cut where the cumulated DRAM read first covers 90 % of one pass
    cum = 0.0
    for i in range(1, len(allm)):
        cum += to_bytes(*allm[i]["dram__bytes_read.sum"])
        if cum >= 0.9 * 10e6 * 384 * 2:
            allm = allm[:i + 1]
            break
    segs = allm[1:] if len(allm) > 1 else allm
    big = max(segs, key=lambda m: to_ms(*m["gpu__time_duration.sum"]))
    md.append("## K3 main pass, C3 (10M x 384, nq=1024, k=100)\n\n`ncu --set full --clock-control none --import-source on`, "
              "all %d K3 launches of the first search (1 warm start + %d segments). Cold-ish caches, serialised: compare shares, not absolutes.\n" % (len(allm), len(segs)))
    md.append("| launch | duration | DRAM read | DRAM write | tensor pipe active | SM clock |\n|---|---|---|---|---|---|")
    for i, m in enumerate(allm):
        md.append("| %d%s | %s %s | %s %s | %s %s | %s %% | %s %s |" % (
            i, " (warm start)" if i == 0 and len(allm) > 1 else "", *m["gpu__time_duration.sum"], *m["dram__bytes_read.sum"], *m["dram__bytes_write.sum"],
            m.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", ("n/a", ""))[0], *m["sm__cycles_elapsed.avg.per_second"]))
    md.append("\nLargest segment:\n\n| metric | value |\n|---|---|")
    for k, (v, u) in big.items():
        md.append("| `%s` | %s %s |" % (k, v, u))
    tot = sum(to_bytes(*m["dram__bytes_read.sum"]) + to_bytes(*m["dram__bytes_write.sum"]) for m in segs)
    traffic["c3:k3_tf32x3_topk:n1"] = {"bytes_per_launch": tot, "source": "profiles/%s_summary.md (ncu --set full, summed over the %d segment launches of one main pass)" % (R, len(segs))}
    md.append("\nDRAM traffic of the main pass (read + write, all segments) = **%.2f GB**; the rows are 15.36 GB as fp32, 7.68 GB as the bf16 copies the screening reads.\n" % (tot / 1e9))

for tag, rep, wl, kern, note in [
        ("K2 scan, C3 nq=1 (10M x 384, k=100; bf16 copies, 232 candidates, exact fp32 re-rank behind it)", "%s_k2_c3q1.ncu-rep" % R, "c3q1", "scan_l2_topk_kernel", "one scan launch")]:
    rep_path = os.path.join(G, rep)
    if not os.path.exists(rep_path):
        continue
    # two consecutive launches are captured: the scan over the bf16 copies and, right behind it, the
    # conditional fp32 repeat, which exits at once unless the re-rank raised its flag -- keep the real one
    cands = raw_metrics_all(rep_path, WANTED)
    m = max(cands, key=lambda mm: to_ms(*mm["gpu__time_duration.sum"]))
    md.append("## %s\n\n`ncu --set full --clock-control none --import-source on` (%s). Cold-ish caches, serialised: compare shares, not absolutes.\n" % (tag, note))
    md.append("| metric | value |\n|---|---|")
    for k, (v, u) in m.items():
        md.append("| `%s` | %s %s |" % (k, v, u))
    if "dram__bytes_read.sum" in m:
        tot = to_bytes(*m["dram__bytes_read.sum"]) + to_bytes(*m["dram__bytes_write.sum"])
        traffic["%s:%s:n1" % (wl, kern)] = {"bytes_per_launch": tot, "source": "profiles/%s_summary.md (ncu --set full, one launch)" % R}
        md.append("\nDRAM traffic per launch (read + write) = **%.2f GB**." % (tot / 1e9))
    md.append("")

for tag, path in [("bench.py --workload c3 --steps 2 --warmup 1", "%s_c3_launches.csv" % R),
                  ("bench.py --workload c3q1 --steps 5 --warmup 3", "%s_q1_launches.csv" % R)]:
    fp = os.path.join(G, path)
    if not os.path.exists(fp):
        continue
    L = launches(fp)
    with open(os.path.join(P, path), "w") as f:
        f.write("launch,kernel,gpu__time_duration_us\n")
        for i, (k, v) in enumerate(L):
            f.write("%d,%s,%.3f\n" % (i, k, v))
    agg = OrderedDict()
    for k, v in L:
        agg.setdefault(k, []).append(v)
    md.append("## Launch list: `%s` (first %d launches, `--metrics gpu__time_duration.sum --clock-control none`)\n" % (tag, len(L)))
    md.append("Full list: `profiles/%s`.\n" % path)
    md.append("| kernel | launches | mean us | min us | max us | share of listed GPU time |\n|---|---|---|---|---|---|")
    total = sum(v for _, v in L)
    for k, vs in agg.items():
        md.append("| `%s` | %d | %.1f | %.1f | %.1f | %.1f %% |" % (k, len(vs), sum(vs) / len(vs), min(vs), max(vs), 100 * sum(vs) / total))
    md.append("")

open(os.path.join(P, "%s_summary.md" % R), "w").write("\n".join(md) + "\n")
tj = os.path.join(P, "traffic.json")
old = json.load(open(tj)) if os.path.exists(tj) else {}
old.update(traffic)
json.dump(old, open(tj, "w"), indent=1)
print("\n".join(md))
