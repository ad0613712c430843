#!/usr/bin/env python
"""Per-kernel census of the Blackwell-only SASS instructions in libb2flat.so (cuobjdump -sass):
UTCHMMA / UTCQMMA (tcgen05.mma), UTMALDG (TMA loads), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit),
SYNCS (mbarrier), REDUX, plus FFMA / HFMA2 for the CUDA-core kernels.

    python scripts/sass_census.py > profiles/r02_sass_census.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "nv_genai_b200", "lib", "libb2flat.so")
OPS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "SYNCS", "REDUX", "FFMA", "HFMA2", "LDG", "STG"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            name = name.replace("(anonymous namespace)::", "").replace("void ", "")
            depth, cut = 0, len(name)
            for i, ch in enumerate(name):  # template arguments stay, the parameter list goes
                if ch == "<":
                    depth += 1
                elif ch == ">":
                    depth -= 1
                elif ch == "(" and depth == 0:
                    cut = i
                    break
            name = name[:cut]
            per[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    per[name][o] += 1
            per[name]["_total"] += 1
    print("# SASS census of nv_genai_b200/lib/libb2flat.so (sm_100a), `cuobjdump -sass`, instruction counts per kernel")
    print("# " + " ".join(f"{o:>8s}" for o in OPS) + "    total  kernel")
    tot = collections.Counter()
    for k, c in per.items():
        tot.update(c)
        print("  " + " ".join(f"{c[o]:8d}" for o in OPS) + f" {c['_total']:8d}  {k}")
    print("# " + " ".join(f"{tot[o]:8d}" for o in OPS) + f" {tot['_total']:8d}  ALL KERNELS ({len(per)})")


if __name__ == "__main__":
    sys.exit(main())
