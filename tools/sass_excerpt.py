#!/usr/bin/env python
"""cuobjdump -sass excerpt of the innermost tap loop (the 8-tap run body / the steady sliding block) of a stencil kernel:
the smallest backward-branch loop that contains packed fp32x2 instructions, with an opcode histogram in front.
  python tools/sass_excerpt.py <libvip_b200.so> <mode 0|1|2> <lut kind> > profiles/<name>.sass.txt"""
import re
import subprocess
import sys
from collections import Counter

so, mode, lk = sys.argv[1], sys.argv[2], sys.argv[3]
names = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funs = re.findall(r"Function : (\S+)", names)
P = {"0": 8, "1": 8, "2": 4}[mode]
want = [f for f in funs if f"stencil_kernelILi{mode}ELi{P}ELi{lk}ELb0ELb0EE" in f]
assert want, "kernel not found"
out = subprocess.run(["cuobjdump", "-sass", "-fun", want[0], so], capture_output=True, text=True).stdout
lines = [l for l in out.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l)]
addr = lambda l: int(re.match(r"\s+/\*([0-9a-f]{4,5})\*/", l).group(1), 16)
op = lambda l: re.sub(r"^\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d\s+)?", "", l).split()[0]
best = None
for l in lines:
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", l)
    if m and int(m.group(1), 16) < addr(l):
        a0, a1 = int(m.group(1), 16), addr(l)
        body = [x for x in lines if a0 <= addr(x) <= a1]
        if any("FFMA2" in x for x in body) and (best is None or len(body) < len(best)):
            best = body
c = Counter(op(x).split(".")[0] for x in best)
demangled = subprocess.run(["c++filt", want[0]], capture_output=True, text=True).stdout.strip()
print(f"# {demangled}")
print(f"# innermost tap loop: {len(best)} SASS instructions for 64 taps (8 outputs x 8 taps; ABF: 32 taps) of one thread")
print("# opcode histogram: " + ", ".join(f"{k} {v}" for k, v in c.most_common()))
for x in best:
    print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", x))
