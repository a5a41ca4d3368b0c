#!/usr/bin/env python3
"""Summarise ncu artefacts brought back in gpurun_out/ into small text files under profiles/.

    python tools/ncu_summary.py report gpurun_out/prof.ncu-rep profiles/rNN_name_ncu.txt
    python tools/ncu_summary.py launches gpurun_out/launches.csv profiles/rNN_name_launches.txt
"""
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
]


def report(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write("# ncu --set full --clock-control none (cold-cache, serialised replays): %s\n" % rep)
        for n, r in enumerate(rows[2:]):
            f.write("\n[launch %d] %s\n" % (n, r[hdr.index("Kernel Name")][:140]))
            for m in METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    f.write("  %-62s %s %s\n" % (m, r[i], units[i]))
            stall = [(hdr[i], float(r[i])) for i in range(len(hdr))
                     if hdr[i].startswith("smsp__average_warps_issue_stalled") and hdr[i].endswith("_per_issue_active.ratio")
                     and r[i] not in ("", "n/a")]
            for name, v in sorted(stall, key=lambda x: -x[1])[:6]:
                f.write("  stall %-56s %.3f\n" % (name.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))


def launches(path, out):
    rows = [r for r in csv.reader(open(path, errors="ignore")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = {}
    for r in rows:
        if r is hdr or len(r) <= iv:
            continue
        try:
            v = float(r[iv].replace(",", ""))
        except ValueError:
            continue
        k = r[ik].split("(")[0][:90]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none: per-kernel totals (ns; cold-cache, serialised)\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("%6.2f%%  n=%5d  total=%12.0f ns  avg=%10.0f ns  %s\n" % (100 * t / tot, n, t, t / n, k))


if __name__ == "__main__":
    {"report": report, "launches": launches}[sys.argv[1]](sys.argv[2], sys.argv[3])
