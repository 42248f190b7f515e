#!/usr/bin/env python
"""Per-kernel tally of the Blackwell-specific SASS opcodes in libocc_b200.so (B200_PROFILING.md's mnemonic table):
UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld/st), UTMALDG / UTMASTG (TMA load / store), UTCBAR (tcgen05.commit),
HMMA (legacy mma.sync), MUFU.   python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt   (no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "occm_b200", "libocc_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "HMMA", "MUFU", "SYNCS", "BAR.SYNC"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = {}
    names = sorted(set(re.findall(r"Function : (\S+)", sass)))
    if names:
        out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        demangle = dict(zip(names, out))
    cur = None
    tally = collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = demangle.get(m.group(1), m.group(1))
            cur = re.sub(r"\(.*", "", cur)
            tally[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        tally[cur]["_total"] += 1
        for o in OPS:
            if op.startswith(o):
                tally[cur][o] += 1
    print("# SASS opcode tally of occm_b200/libocc_b200.so (cuobjdump -sass, sm_100a); columns:", " ".join(OPS), "total")
    tot = collections.Counter()
    for k, c in tally.items():
        if not any(c[o] for o in OPS if o not in ("MUFU", "BAR.SYNC", "SYNCS")) and "kernel" not in k:
            continue
        print(f"{k[:90]:<90} " + " ".join(f"{c[o]:>7d}" for o in OPS) + f" {c['_total']:>7d}")
        tot.update(c)
    print(f"{'TOTAL':<90} " + " ".join(f"{tot[o]:>7d}" for o in OPS) + f" {tot['_total']:>7d}")


if __name__ == "__main__":
    main()
