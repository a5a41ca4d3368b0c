#!/usr/bin/env python3
"""Top stalled SASS instructions of one kernel in an ncu report (--import-source on):
    python tools/ncu_sass.py rep.ncu-rep [kernel-index] [top-N] [lo hi]   (lo..hi: dump that SASS row range)"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
k, end = starts[which], starts[which + 1]
hdr = rows[k + 1]
isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
out = [(n, r[isrc].strip(), int(r[iex] or 0), int(r[ismp] or 0)) for n, r in enumerate(rows[k + 2:end])
       if len(r) > ismp and r[iex].isdigit()]
tot = sum(o[3] for o in out) or 1
print(rows[k][1][:80], "| samples", tot, "| warp-inst", sum(o[2] for o in out))
if len(sys.argv) > 5:
    lo, hi = int(sys.argv[4]), int(sys.argv[5])
    for n, s, e, sm in out:
        if lo <= n <= hi:
            print("%4d %9d %6d  %s" % (n, e, sm, s))
else:
    for n, s, e, sm in sorted(out, key=lambda o: -o[3])[:top]:
        print("%4d %9d %6d %5.1f%%  %s" % (n, e, sm, 100.0 * sm / tot, s))
