#!/usr/bin/env python
"""Turns ncu outputs brought back in gpurun_out/ into the small tracked summaries under profiles/.

    python tools_profile_summary.py launches gpurun_out/r01_launches.csv profiles/r01_launch_summary.md "<command>"
    python tools_profile_summary.py kernels  gpurun_out/x.ncu-rep profiles/x.md "<command>" [regex]
"""
import csv
import io
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__pcsamp_warps_issue_stalled_long_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_short_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait", "smsp__pcsamp_sample_count"]


def short(name):
    m = re.search(r"(k_[a-z_]+)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:40]


def launches(src, dst, cmd):
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    seq = [(short(r[ik]), float(r[iv].replace(",", "")) / 1e3) for r in rows[1:]]
    # the last full scan step: from the last k_scan launch to the end
    last = max(i for i, (k, _) in enumerate(seq) if k.startswith("k_scan"))
    step = seq[last:]
    prev = [i for i, (k, _) in enumerate(seq[:last]) if k.startswith("k_scan")]
    if prev:   # kernels of the same step launched on side streams may be listed before k_scan's row of the next step
        step = seq[prev[-1]:last]
    total = sum(t for _, t in step)
    with open(dst, "w") as f:
        f.write("# Launch list summary (one scan step, one batch at a time)\n\n`%s`\n\n" % cmd)
        f.write("ncu serialises the launches and flushes caches between them: compare shares, not absolutes. In the pipeline\n"
                "k_prefix_*/k_pods run beside k_confirm/k_verify, the k_sort_* kernels beside k_plan/k_check, k_check_long beside k_check.\n\n")
        f.write("| kernel | us | share of the step |\n|---|---|---|\n")
        for k, t in step:
            f.write("| %s | %.1f | %.3f |\n" % (k, t, t / total))
        f.write("| **sum** | %.1f | 1.000 |\n" % total)
    print(open(dst).read())


def kernels(src, dst, cmd, regex=None):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write("# ncu --set full summary\n\n`%s`\n\n" % cmd)
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))
            if regex and not re.search(regex, d["Kernel Name"]):
                continue
            f.write("## %s\n\n| metric | value |\n|---|---|\n" % short(d["Kernel Name"]))
            for k in KEYS:
                if d.get(k) not in (None, ""):
                    f.write("| %s | %s %s |\n" % (k, d[k], u.get(k, "")))
            f.write("\n")
    print(open(dst).read())


if __name__ == "__main__":
    {"launches": launches, "kernels": kernels}[sys.argv[1]](*sys.argv[2:])
