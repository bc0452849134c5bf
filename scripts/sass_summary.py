#!/usr/bin/env python
"""Opcode histogram of the shipped library's SASS, per kernel: evidence that the GEMMs are tcgen05 / TMEM / TMA code
(UTCHMMA = tcgen05.mma, UTMALDG = cp.async.bulk.tensor, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit,
SYNCS = mbarrier ops) and that nothing multicasts.  Needs cuobjdump (CUDA toolkit); no GPU.

    python scripts/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "multidisttorch_b200", "csrc", "libmdt_b200.so")
KEY = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCATOMSWS", "SYNCS", "HMMA", "FFMA", "MUFU",
       "LDG", "STG", "LDS", "STS", "ATOM", "RED", "BAR", "ACQBULK", "UCGABAR_ARV"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = per.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
            if "MULTICAST" in m.group(2):
                cur["*MULTICAST*"] += 1
    dm = demangle(list(per))
    total = collections.Counter()
    print(f"# SASS opcode histogram of {os.path.relpath(LIB, ROOT)} (sm_100a), cuobjdump -sass; columns = instruction counts")
    print("# UTCHMMA=tcgen05.mma  UTMALDG=TMA tensor load  LDTM/STTM=tcgen05.ld/st  UTCBAR=tcgen05.commit  SYNCS=mbarrier  *MULTICAST* must be 0")
    for fn, c in per.items():
        name = re.sub(r"\(.*", "", dm.get(fn, fn)).replace("void ", "")
        keys = [k for k in KEY if c.get(k)]
        print(f"{name:<58s} insts={sum(c.values()):6d}  " + "  ".join(f"{k}={c[k]}" for k in keys) +
              (f"  *MULTICAST*={c['*MULTICAST*']}" if c.get("*MULTICAST*") else ""))
        total.update(c)
    print("# TOTAL  " + "  ".join(f"{k}={total[k]}" for k in KEY if total.get(k)) + f"  *MULTICAST*={total.get('*MULTICAST*', 0)}")


if __name__ == "__main__":
    sys.exit(main())
