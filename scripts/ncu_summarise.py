#!/usr/bin/env python
"""Turn ncu output into the small text summaries kept under profiles/.

  python scripts/ncu_summarise.py launches gpurun_out/launches.csv profiles/r01_x_launches.txt
      per-kernel launch count / total / mean duration and share of the profiled command
  python scripts/ncu_summarise.py full gpurun_out/prof.ncu-rep profiles/r01_x_full.txt
      the headline metrics of every captured launch (`ncu --set full`), via `ncu -i --page raw --csv`
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__t_sectors.sum",
    "l1tex__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
]


def launches(src, dst):
    text = open(src).read()
    start = text.find('"ID"')
    rows = list(csv.reader(io.StringIO(text[start:])))
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv:
            continue
        name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").strip()
        v = float(r[iv].replace(",", ""))
        u = r[iu]
        us = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (serialised, cold-cache per-launch times)\n")
        f.write("# source: %s ; total %.1f us over %d launches\n" % (src, total, sum(a[0] for a in agg.values())))
        f.write("%-90s %8s %12s %10s %7s\n" % ("kernel", "launches", "total_us", "mean_us", "share"))
        for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%-90s %8d %12.1f %10.2f %6.1f%%\n" % (name[:90], a[0], a[1], a[1] / a[0], 100 * a[1] / total))
    print(open(dst).read())


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on ; source report: %s\n" % src)
        for r in rows[2:]:
            f.write("\n== %s ==\n" % r[hdr.index("Kernel Name")][:150])
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write("%-88s %18s %s\n" % (k, r[i], units[i]))
    print(open(dst).read())


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
