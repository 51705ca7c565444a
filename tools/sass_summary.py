#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of libxgnn_b200.so (cuobjdump -sass): evidence of what the hot kernels are made of.
  python tools/sass_summary.py > profiles/sass_r02.txt"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "xgnn_b200" / "lib" / "libxgnn_b200.so"
WATCH = ["LDGSTS", "LDG", "STG", "LDS", "STS", "ATOMG", "REDG", "RED", "ATOM", "LDGDEPBAR", "DEPBAR", "BAR", "SHFL", "VOTE",
         "MATCH", "IMAD", "UBLKCP", "SYNCS", "UTMALDG", "UTCMMA", "HMMA", "NANOSLEEP", "MEMBAR"]
out = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
kernels = OrderedDict()
cur = None
arch = set()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = kernels.setdefault(m.group(1), Counter())
        continue
    m = re.search(r"arch = (sm_\w+)", ln)
    if m:
        arch.add(m.group(1))
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", ln)
    if m and cur is not None:
        cur[m.group(1)] += 1
        cur["_total"] += 1
        full = m.group(1) + m.group(2)
        if m.group(1) in ("LDG", "STG", "ATOMG", "REDG", "LDGSTS", "ATOM", "RED"):
            cur["~" + full] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
print(f"# cuobjdump -sass {LIB.relative_to(ROOT)}   cubins: {sorted(arch)}")
print("# no UTCMMA / HMMA is expected: nothing on this path is a contraction; rows are staged with cp.async (LDGSTS) in the ring "
      "gather and with TMA bulk copies (UBLKCP + mbarrier SYNCS) in gather_rows_bulk_kernel")
for (name, c), dn in zip(kernels.items(), demangle):
    short = dn.replace("(anonymous namespace)::", "").replace("void ", "")
    short = re.sub(r"\(.*", "", short)[:110]
    if c["_total"] < 40:
        continue
    print(f"\n{short}   [{c['_total']} instructions]")
    print("   " + "  ".join(f"{k}={c[k]}" for k in WATCH if c[k]))
    wide = {k[1:]: v for k, v in c.items() if k.startswith("~")}
    print("   " + "  ".join(f"{k}={v}" for k, v in sorted(wide.items())))
