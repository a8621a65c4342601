#!/usr/bin/env python
"""Summarise ncu outputs into markdown for profiles/.

  summarize_ncu.py launches <launches.csv>          per-kernel time totals / shares
  summarize_ncu.py full <report.ncu-rep> [regex]    key metrics of each captured launch
"""
import csv, re, subprocess, sys, collections

def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "")
    return name[:70]

def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]; ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[1:]:
        if r[mi] != "gpu__time_duration.sum": continue
        k = short(r[ki]); tot[k] += float(r[vi].replace(",", "")); cnt[k] += 1
    T = sum(tot.values())
    print("| kernel | launches | total us | share |\n|---|---|---|---|")
    for k, v in tot.most_common(14):
        print("| `%s` | %d | %.1f | %.1f %% |" % (k, cnt[k], v / 1e3, 100 * v / T))
    print("\ntotal %.1f us over %d launches (cold-cache, serialised: shares only)" % (T / 1e3, sum(cnt.values())))

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct",
        "smsp__warp_issue_stalled_wait_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_not_selected_per_warp_active.pct",
        "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_dispatch_stall_per_warp_active.pct"]

def full(path, pat=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        if pat and not re.search(pat, r[ki]): continue
        print("### `%s`  grid %s block %s\n" % (short(r[ki]), r[hdr.index("Grid Size")], r[hdr.index("Block Size")]))
        print("| metric | value | unit |\n|---|---|---|")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w); print("| %s | %s | %s |" % (w, r[i], units[i]))
        print()

if __name__ == "__main__":
    if sys.argv[1] == "launches": launches(sys.argv[2])
    else: full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
