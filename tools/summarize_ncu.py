"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches.csv profiles/r01_launches.txt
  python tools/summarize_ncu.py full gpurun_out/prof_x.ncu-rep profiles/r01_prof_x.txt
"""
import collections
import csv
import re
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg.per_second", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_membar_per_warp_active.pct",
]


def short(name):
    name = re.sub(r"^.*?unnamed>::", "", name)
    return re.sub(r"\(.*$", "", name)


def launches(src, dst):
    rows = list(csv.DictReader(l for l in open(src) if l.startswith('"')))
    per = collections.OrderedDict()
    total = 0.0
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        k = short(r["Kernel Name"])
        c = per.setdefault(k, [0, 0.0])
        c[0] += 1
        c[1] += ns
        total += ns
    with open(dst, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none ; source: %s\n" % src)
        f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
        f.write("%-34s %8s %12s %8s\n" % ("kernel", "launches", "total_us", "share"))
        for k, (n, ns) in sorted(per.items(), key=lambda kv: -kv[1][1]):
            f.write("%-34s %8d %12.1f %7.2f%%\n" % (k, n, ns / 1e3, 100 * ns / total))
        f.write("%-34s %8d %12.1f\n" % ("TOTAL", sum(v[0] for v in per.values()), total / 1e3))


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [i for i, h in enumerate(hdr) if h in KEEP]
    name_col = hdr.index("Kernel Name")
    grid_col = hdr.index("Grid Size") if "Grid Size" in hdr else None
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on ; source: %s\n" % src)
        for n, r in enumerate(rows[2:]):
            f.write("launch %d: %s grid %s\n" % (n, short(r[name_col]), r[grid_col] if grid_col is not None else "?"))
            for i in cols:
                f.write("  %-72s %16s %s\n" % (hdr[i], r[i], units[i]))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
