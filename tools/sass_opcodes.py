#!/usr/bin/env python
"""Per-kernel counts of the SASS opcodes that prove which hardware paths libfdb200.so uses:
tcgen05 MMA (UTCHMMA), TMEM loads/stores (LDTM/STTM), TMA bulk copies (UBLKCP), tcgen05 commit (UTCBAR),
mbarrier (SYNCS), SFU (MUFU.*), plus FFMA/LDS/spill markers.

    python tools/sass_opcodes.py [lib] > profiles/r02/sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "fraud_detection_b200", "libfdb200.so")
OPS = ["UTCHMMA", "UTCQMMA", "HMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTCBAR", "SYNCS", "MUFU.EX2", "MUFU.RCP", "MUFU.LG2",
       "FFMA", "FFMA2", "FMUL", "FADD", "DFMA", "LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL", "REDUX", "ATOM", "RED"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
counts, order, cur = {}, [], None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", cur)
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        order.append(cur)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                counts[cur][o] += 1
print(f"# SASS opcode counts per kernel of {os.path.relpath(lib, ROOT)} (cuobjdump -sass; static counts, not executed counts)")
total = collections.Counter()
for k in order:
    c = counts[k]
    total.update(c)
    body = "  ".join(f"{o}={c[o]}" for o in OPS if c[o])
    print(f"\n{k}\n    instructions={c['_total']}  {body}")
print("\nTOTAL\n    " + "  ".join(f"{o}={total[o]}" for o in OPS if total[o]))
