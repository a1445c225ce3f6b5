#!/usr/bin/env python
"""profiles/summarize.py <tag> <launches.csv> <report.ncu-rep> [bench.json] — writes profiles/<tag>_summary.md
from an ncu launch list (--metrics gpu__time_duration.sum) and one --set full capture of render_kernel."""
import collections
import csv
import json
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"]


def launch_table(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except Exception:
            continue
        ms = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui].replace("second", "") or "s", 1e-6) * v
        agg[r[ki][:80]][0] += 1
        agg[r[ki][:80]][1] += ms
    tot = sum(v[1] for v in agg.values())
    out = ["| kernel | launches | total ms | share |", "|---|---|---|---|"]
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:8]:
        out.append(f"| `{k}` | {v[0]} | {v[1]:.3f} | {100 * v[1] / tot:.1f}% |")
    return out


def raw_metrics(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    out = ["| metric | value |", "|---|---|"]
    for w in WANT:
        if w in d:
            out.append(f"| {w} | {d[w][0]} {d[w][1]} |")
    stalls = sorted(((float(v[0]), h) for h, v in d.items()
                     if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")
                     or h.startswith("smsp__average_warp_latency_issue_stalled") and h.endswith(".ratio")), reverse=True)[:6]
    for v, h in stalls:
        out.append(f"| {h} | {v:.3f} |")
    return out, d


if __name__ == "__main__":
    tag, launches, rep = sys.argv[1:4]
    lines = [f"# {tag}", ""]
    if len(sys.argv) > 4:
        b = json.loads(open(sys.argv[4]).read().strip().splitlines()[-1])
        lines += ["## bench.py line (not under ncu)", "", "```json", json.dumps({k: b[k] for k in ("value", "ms_per_step", "e2e", "roofline", "clocks", "gpu_launches", "cpu_baseline", "latency") if k in b}, indent=1), "```", ""]
    lines += ["## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: shares only)", ""] + launch_table(launches) + [""]
    m, d = raw_metrics(rep)
    lines += ["## render_kernel, `ncu --set full --clock-control none --import-source on`", ""] + m + [""]
    open(f"profiles/{tag}_summary.md", "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))
