#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into small tracked text files under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches_r1.csv profiles/r1_launches.md
  python tools/ncu_summary.py full gpurun_out/prof_pair_r1.ncu-rep profiles/r1_pair_full.md
"""
import collections
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_static",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
]


def launches(src, dst):
    rows = list(csv.reader(open(src, errors="replace")))
    hdr = None
    per = collections.OrderedDict()
    for r in rows:
        if len(r) > 5 and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            rec = dict(zip(hdr, r))
            if rec.get("Metric Name") == "gpu__time_duration.sum":
                name = rec["Kernel Name"].split("(")[0].replace("void ", "")
                per.setdefault(name, []).append(float(rec["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in per.values())
    with open(dst, "w") as f:
        f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare SHARES)\n\n")
        f.write("source: %s\n\n| kernel | launches | avg ns | total us | share |\n|---|---|---|---|---|\n" % src)
        for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            f.write("| `%s` | %d | %.0f | %.1f | %.3f |\n" % (k, len(v), sum(v) / len(v), sum(v) / 1e3, sum(v) / tot))
        f.write("\ntotal %.1f us over %d launches\n" % (tot / 1e3, sum(len(v) for v in per.values())))


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none (per launch; caches flushed between replays)\n\nsource: %s\n\n" % src)
        for n, r in enumerate(rows[2:]):
            rec = dict(zip(hdr, r))
            f.write("## launch %d: `%s`\n\n| metric | value | unit |\n|---|---|---|\n" % (n, rec.get("Kernel Name", "?").split("(")[0]))
            for w in WANT:
                if w in rec:
                    f.write("| %s | %s | %s |\n" % (w, rec[w], units[hdr.index(w)]))
            f.write("\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
