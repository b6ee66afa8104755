#!/usr/bin/env python
"""Per-kernel counts of the SASS opcodes that prove tcgen05 / TMA / TMEM use (profiles/r02_sass_opcodes.txt).

    python tools/sass_opcodes.py [lib.so] > profiles/r02_sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "simclr_b200", "libntxent_b200.so")
WANT = re.compile(r"^(UTCHMMA|UTMALDG|LDTM|STTM|UTCBAR|SYNCS|MUFU\.EX2|FFMA2|MULTIMEM|UTMAPF|UTCCP)")

sass = subprocess.run(["cuobjdump", "-sass", lib], check=True, capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                       text=True).stdout.split("\n")
counts, cur, it = collections.OrderedDict(), None, iter(names)
for line in sass.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        full = next(it)
        short = re.sub(r"\(anonymous namespace\)::", "", full)
        depth, out = 0, []
        for ch in short:                      # cut at the argument list, keep template arguments
            if ch == "<":
                depth += 1
            elif ch == ">":
                depth -= 1
            elif ch == "(" and depth == 0:
                break
            out.append(ch)
        cur = "".join(out).replace("void ", "")
        counts[cur] = collections.Counter()
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if m and cur and WANT.match(m.group(1)):
        counts[cur][m.group(1)] += 1

print("# SASS opcode counts of simclr_b200/libntxent_b200.so (cuobjdump -sass, sm_100a), per kernel; tools/sass_opcodes.py.")
print("# UTCHMMA = tcgen05.mma, UTMALDG = TMA tile load (cp.async.bulk.tensor), LDTM / STTM = tcgen05.ld / tcgen05.st (TMEM),")
print("# UTCBAR = tcgen05.commit, SYNCS.* = mbarrier operations, MUFU.EX2 = ex2.approx, FFMA2 = packed fp32x2 FMA,")
print("# MULTIMEM = multimem.st through the NVSwitch multicast mapping.  Template arguments of the tensor-core kernels:")
print("# tc_fwd_kernel<NDB, kScaled, kSplit>, tc_fwd_sym_kernel<NDB, kScaled>, tc_bwd_kernel<NDB, kGate, kScaled, kSplit>.")
print()
for k in sorted(counts):
    if counts[k]:
        print(k)
        print("    " + "  ".join(f"{op}={n}" for op, n in sorted(counts[k].items(), key=lambda kv: (-kv[1], kv[0]))))
