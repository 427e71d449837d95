#!/usr/bin/env python3
"""Per kernel of liblinalg_b200.so: how often the Blackwell-specific SASS mnemonics occur (cuobjdump -sass; no GPU needed).
   python tools/sass_counts.py > profiles/r2_sass_counts.txt
UTCHMMA / UTCIMMA = tcgen05.mma (f16/tf32 kinds / i8 kind), .2CTA = cta_group::2; LDTM = tcgen05.ld; UTMALDG / UTMASTG = TMA tensor
load / store; UBLKCP = cp.async.bulk; SYNCS = mbarrier; FFMA2 = packed f32x2 FMA; DFMA / IMAD for the f64 / integer SIMT engines."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "linear_algebra_b200", "liblinalg_b200.so")
PATTERNS = ["UTCHMMA", "UTCIMMA", "2CTA", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "FFMA2", "FFMA", "DFMA", "IMAD", "LDS.128", "STS.128", "LDG.E.128", "STG.E.128", "SHFL", "REDUX", "ATOM", "RED"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    name = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            name = re.sub(r"\(.*", "", name)                  # drop the parameter list
            name = re.sub(r"^void ", "", name)
            counts[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for p in PATTERNS:
            if p == "FFMA" and op.startswith("FFMA2"):
                continue
            if p == "2CTA":
                if ".2CTA" in op:
                    counts[name][p] += 1
            elif p == "RED":
                if op.startswith("RED.") or op == "RED":
                    counts[name][p] += 1
            elif op.startswith(p):
                counts[name][p] += 1
    print("# SASS mnemonic counts per kernel of linear_algebra_b200/liblinalg_b200.so (tools/sass_counts.py, cuobjdump -sass, sm_100a)")
    print("# columns: " + " ".join(PATTERNS))
    tot = collections.Counter()
    for name, c in counts.items():
        tot.update(c)
        if not any(c[p] for p in ("UTCHMMA", "UTCIMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "FFMA2", "DFMA")) and len(sys.argv) < 2:
            continue                                         # plain kernels are listed only with an argument
        print("%-110s %s" % (name[:110], " ".join("%s=%d" % (p, c[p]) for p in PATTERNS if c[p])))
    print("\n# whole library (%d kernels): %s" % (len(counts), " ".join("%s=%d" % (p, tot[p]) for p in PATTERNS if tot[p])))


if __name__ == "__main__":
    main()
