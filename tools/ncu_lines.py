#!/usr/bin/env python3
"""Per-CUDA-source-line summary of an ncu report captured with --import-source on (needs -lineinfo):
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [kernel-index] [top-N]
prints, for the chosen launch, warp-instructions executed and stall samples per source line."""
import csv
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
# split into kernels at "File Path" rows that follow a blank/first line; the header row starts with "Line No"
kern, cur = [], None
for r in rows:
    if r and r[0] == "Line No":
        cur = {"hdr": r, "rows": []}
        kern.append(cur)
    elif cur is not None and r and r[0].isdigit():
        cur["rows"].append(r)
k = kern[which]
h = k["hdr"]
iex, ismp, isrc = h.index("Instructions Executed"), h.index("# Samples"), 1
iaddr = h.index("Address")
lines = [(int(r[0]), r[isrc], int(r[iex] or 0), int(r[ismp] or 0)) for r in k["rows"] if r[iaddr] == "-"]
tot_i = sum(l[2] for l in lines) or 1
tot_s = sum(l[3] for l in lines) or 1
print("launch %d: %d warp-instructions, %d samples" % (which, tot_i, tot_s))
print("by instructions:")
for ln, src, i, s in sorted(lines, key=lambda l: -l[2])[:top]:
    print("%5d %6.2f%% inst %6.2f%% smp  %s" % (ln, 100.0 * i / tot_i, 100.0 * s / tot_s, src.strip()[:110]))
print("by samples:")
for ln, src, i, s in sorted(lines, key=lambda l: -l[3])[:top // 2]:
    print("%5d %6.2f%% inst %6.2f%% smp  %s" % (ln, 100.0 * i / tot_i, 100.0 * s / tot_s, src.strip()[:110]))
