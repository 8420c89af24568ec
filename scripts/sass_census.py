#!/usr/bin/env python
"""SASS opcode census of libomnia_recall.so: which kernels carry tcgen05 / TMA / mbarrier instructions.

    python scripts/sass_census.py > profiles/r02_sass_opcodes.txt

Runs `cuobjdump -sass` (no GPU needed) and `c++filt`, counts the opcodes that matter per kernel."""
import collections
import re
import subprocess
import sys
from pathlib import Path

LIB = Path(__file__).resolve().parent.parent / "omnia_b200" / "lib" / "libomnia_recall.so"
WANT = ("UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "UBLKCP", "LDTM", "UTCATOMSWS", "SYNCS", "REDUX", "CREDUX", "ATOMG", "ATOMS",
        "ATOM", "RED", "DFMA", "ELECT", "R2UR", "ACQBULK", "PREEXIT", "MATCH")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            op = m.group(1)
            kernels[cur]["instructions"] += 1
            base = op.split(".")[0]
            if base in WANT:
                key = base + (".2CTA" if ".2CTA" in op else "")
                kernels[cur][key] += 1
    names = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    print("# SASS opcode counts per kernel of libomnia_recall.so (cuobjdump -sass, sm_100a); scripts/sass_census.py.")
    print("# UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk,")
    print("# UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, ACQBULK / PREEXIT = griddepcontrol.wait / launch_dependents,")
    print("# ELECT / R2UR = elect.sync and vector-to-uniform register moves (the MMA issue loop has 1 ELECT and 6 R2UR per")
    print("# K block; before the operands were made warp-uniform it had 6 ELECT and 22 R2UR.BROADCAST).")
    print()
    for (mangled, cnt), name in sorted(zip(kernels.items(), names), key=lambda t: t[1]):
        name = re.sub(r"omr::\(anonymous namespace\)::", "", name)
        name = re.sub(r"^void ", "", name)
        rest = "  ".join(f"{k}={v}" for k, v in sorted(cnt.items()) if k != "instructions")
        print(name[:150])
        print(f"    instructions {cnt['instructions']}  {rest}")


if __name__ == "__main__":
    sys.exit(main())
