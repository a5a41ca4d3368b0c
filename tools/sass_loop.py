"""Instruction mix of the streaming loop of a kernel in libb200recon.so: the shortest backward branch whose body holds
LDGSTS (the software pipeline of chunk_pipeline<..., SAFE_Z=true> in integrate_items_kernel).
    python tools/sass_loop.py [lib.so] [mangled-name substring] [which loop, by length: 0 = shortest] [dump]"""
import re, subprocess, sys, collections
lib = sys.argv[1] if len(sys.argv) > 1 else "3d-reconstruction-planning_b200/lib/libb200recon.so"
pat = sys.argv[2] if len(sys.argv) > 2 else "integrate_items_kernelILi1ELb1ELb0ELi1024"
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, ins = None, []
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    if fn and pat in fn:
        m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: i for i, (a, _) in enumerate(ins)}
loops = []
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a and int(m.group(1), 16) in addr:
        j = addr[int(m.group(1), 16)]
        if any("LDGSTS" in x for _, x in ins[j:i + 1]):
            loops.append((j, i))
loops.sort(key=lambda l: l[1] - l[0])
print("kernel instructions: %d; loops with LDGSTS (start, end, length): %s" % (len(ins), [(hex(ins[j][0]), hex(ins[i][0]), i - j + 1) for j, i in loops]))
j, i = loops[which]
mix = collections.Counter()
for _, t in ins[j:i + 1]:
    t = re.sub(r"^@!?U?P\w+\s+", "", t)
    mix[t.split()[0].split(".")[0]] += 1
print("loop %s..%s: %d instructions" % (hex(ins[j][0]), hex(ins[i][0]), i - j + 1))
print("  " + ", ".join("%s %d" % kv for kv in mix.most_common()))
if len(sys.argv) > 4:
    for a, t in ins[j:i + 1]:
        print("  /*%04x*/ %s" % (a, t))
