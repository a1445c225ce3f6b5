#!/usr/bin/env python
"""profiles/sass_histogram.py [lib.so] — opcode histogram of the shipped library's SASS (cuobjdump -sass), whole library
and per kernel for the render kernels; writes profiles/r02_sass_histogram.txt.  Lets a reader check the TMA / mbarrier /
packed-FP32x2 claims of DESIGN.md section 3.3 without rebuilding."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "libzl_b200/libzl_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
ins = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z][A-Z0-9_]*)")
per = collections.OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        per[cur] = collections.Counter()
        continue
    m = ins.match(line)
    if m and cur:
        per[cur][m.group(1)] += 1
demangled = dict(zip(per, subprocess.run(["c++filt"], input="\n".join(per), capture_output=True, text=True).stdout.splitlines()))
total = collections.Counter()
for c in per.values():
    total.update(c)
out = [f"# SASS opcode histogram of {lib} (cuobjdump -sass, sm_100a); {len(per)} kernels, {sum(total.values())} instructions", ""]
mark = ["UBLKCP", "SYNCS", "UTMALDG", "UTMASTG", "LDGSTS", "FMUL2", "FADD2", "FFMA2", "CREDUX", "REDUX", "DFMA", "DADD", "F2F", "I2F", "F2I",
        "HMMA", "UTCHMMA", "LDS", "STS", "LDG", "STG", "SHFL"]
out.append("## Marker opcodes, whole library")
out += [f"{k:10s} {total.get(k, 0)}" for k in mark]
out += ["", "## Whole library, top 60"]
out += [f"{n:7d} {k}" for k, n in total.most_common(60)]
for name, c in per.items():
    d = demangled.get(name, name)
    if "gather_kernel" in d or "plan_kernel" in d or "walk_kernel" in d or "master_mix" in d or "master_pcm" in d:
        out += ["", f"## {d[:200]}", f"instructions {sum(c.values())}; " + ", ".join(f"{k} {c[k]}" for k in mark if c.get(k))]
open("profiles/r02_sass_histogram.txt", "w").write("\n".join(out) + "\n")
print("\n".join(out[:30]))
