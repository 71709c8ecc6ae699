#!/usr/bin/env python
"""Opcode histogram of every kernel in band_b200/lib/libbandcu.so (cuobjdump -sass), here, without a GPU.
Shows which Blackwell / Hopper+ machinery each kernel family really uses: UBLKCP (cp.async.bulk, the TMA engine in 1-D
mode), SYNCS (mbarrier), LDGSTS (cp.async), VIADDMNMX / VIMNMX .16x2 (DPX packed min/max), IDP (dp4a), PRMT; ST = ST + STG, LDG = LD + LDG; and that none of
them contains tensor-core (UTCMMA / HMMA / IMMA) or tensor-map (UTMALDG) instructions: the path has no contraction.

  python scripts/sass_histogram.py > profiles/rNN_sass_opcodes.txt
"""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "band_b200", "lib", "libbandcu.so")
KEY = ["UBLKCP", "SYNCS", "LDGSTS", "VIADDMNMX", "VIMNMX", "IDP", "PRMT", "LDS", "STS", "LDG", "ST", "IMAD", "SHFL", "BAR",
       "UTMALDG", "UTCMMA", "HMMA", "IMMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = OrderedDict()
    cur = None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = kernels.setdefault(m.group(1), Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m and cur is not None:
            op = m.group(1)
            cur["ST" if op in ("ST", "STG") else "LDG" if op in ("LD", "LDG") else op] += 1
    demangle = subprocess.run(["c++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    names = dict(zip(kernels, demangle)) if len(demangle) == len(kernels) else {k: k for k in kernels}
    print("# %s: %d kernels, sm_100a SASS (static instruction counts per kernel)" % (os.path.relpath(LIB, ROOT), len(kernels)))
    print("# " + " ".join("%s" % k for k in KEY) + " | total")
    fam = OrderedDict()
    for k, c in kernels.items():
        n = names[k].replace("(anonymous namespace)::", "").replace("bandcu::", "")
        n = re.sub(r"^void ", "", re.sub(r"\(.*", "", n))
        print("%-96s %s | %d" % (n[:96], " ".join("%d" % c.get(x, 0) for x in KEY), sum(c.values())))
        f = re.sub(r"<.*", "", n)
        agg = fam.setdefault(f, Counter())
        for x in KEY:
            agg[x] += c.get(x, 0)
        agg["_n"] += 1
    print("\n# per kernel family (sum over its template instances)")
    print("# %-40s %s" % ("family (instances)", " ".join("%9s" % k for k in KEY)))
    for f, c in fam.items():
        print("  %-40s %s" % ("%s (%d)" % (f, c["_n"]), " ".join("%9d" % c.get(x, 0) for x in KEY)))
    tot = Counter()
    for c in fam.values():
        tot.update({x: c.get(x, 0) for x in KEY})
    print("\n# library totals: " + ", ".join("%s %d" % (x, tot[x]) for x in KEY))
    assert tot["UTCMMA"] == 0 and tot["HMMA"] == 0 and tot["IMMA"] == 0, "tensor-core instructions found"


if __name__ == "__main__":
    sys.exit(main())
