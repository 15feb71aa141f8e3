#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove which hardware paths librgbdfe.so uses (tcgen05: UTCIMMA / UTCBAR,
tensor memory: LDTM, TMA: UTMALDG, legacy mma.sync: IMMA, packed integer work: IDP, VIMNMX, POPC; FP64: DFMA / DADD / DMUL).
Writes profiles/r2_sass_grep.txt.  Runs on the CPU container (cuobjdump only)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "rgbd-slam_b200", "lib", "librgbdfe.so")
PAT = ["UTCIMMA", "UTCBAR", "LDTM", "UTMALDG", "SYNCS", "IMMA", "IDP", "VIMNMX", "POPC", "DFMA", "DADD", "DMUL", "FFMA", "FADD", "FMUL", "MUFU", "BAR.SYNC", "ATOMS", "SHFL"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    counts, total = collections.OrderedDict(), collections.Counter()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", name).replace("rgbdfe::", "").replace("(anonymous namespace)::", "")
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            total[cur] += 1
            for p in PAT:
                if op.startswith(p):
                    counts[cur][p] += 1
    out = ["# cuobjdump -sass rgbd-slam_b200/lib/librgbdfe.so (sm_100a), instruction counts per kernel; written by tools/sass_grep.py",
           "# kernel | total | " + " ".join(PAT)]
    for k, c in counts.items():
        out.append(f"{k} | {total[k]} | " + " ".join(f"{p}={c[p]}" for p in PAT if c[p]))
    path = os.path.join(ROOT, "profiles", "r2_sass_grep.txt")
    open(path, "w").write("\n".join(out) + "\n")
    print(path, len(counts), "kernels")


if __name__ == "__main__":
    sys.exit(main())
