#!/usr/bin/env python3
"""Decode the scheduling control bits (stall, write/read scoreboard, wait mask) of a kernel's SASS:
    nvdisasm -hex file.cubin > all.sass ; python tools/sass_ctrl.py all.sass <function-substring> [filter-regex]"""
import re, sys
lines = open(sys.argv[1]).read().splitlines()
fn = sys.argv[2]
flt = re.compile(sys.argv[3]) if len(sys.argv) > 3 else None
active = False
i = 0
while i < len(lines):
    l = lines[i]
    if l.startswith(".text."):
        active = fn in l
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]{16}) \*/', l)
    if active and m and i + 1 < len(lines):
        m2 = re.match(r'\s*/\* (0x[0-9a-f]{16}) \*/', lines[i + 1])
        if m2:
            hi = int(m2.group(1), 16)
            ctrl = (hi >> 41) & 0x1fffff
            stall, wr, rd, wait = ctrl & 0xf, (ctrl >> 5) & 7, (ctrl >> 8) & 7, (ctrl >> 11) & 0x3f
            ins = m.group(2).strip()
            if flt is None or flt.search(ins) or wait:
                print("%s wr=%s rd=%s wait=%s st=%2d  %s" % (m.group(1), wr if wr != 7 else '-', rd if rd != 7 else '-',
                                                             format(wait, '06b'), stall, ins[:90]))
            i += 2
            continue
    i += 1
