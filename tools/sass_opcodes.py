#!/usr/bin/env python
"""Per-kernel counts of the SASS opcodes that prove which engines the library uses (cuobjdump -sass of the built
libpcc_b200.so): UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UTCCP,
UBLKCP (cp.async.bulk), UTMALDG / UTMASTG (TMA tensor copies), LDGSTS (cp.async), SYNCS (mbarrier), REDUX / MATCH / VOTE
(warp primitives), ATOMS / ATOMG / RED.  Usage: python tools/sass_opcodes.py > profiles/sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "demo-learned-point-cloud-compression_b200", "b200pcc", "libpcc_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTCCP", "UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "SYNCS",
       "HMMA", "MATCH", "VOTE", "REDUX", "SHFL", "ATOMS", "ATOMG", "RED", "LDG", "STG", "LDS", "STS", "POPC"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    dem = {}
    names = re.findall(r"Function : (\S+)", txt)
    if names:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        dem = dict(zip(names, out))
    counts = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = dem.get(m.group(1), m.group(1))
            cur = re.sub(r"^void ", "", cur)
            cur = re.sub(r"\((?!bool\)|int\)|unsigned)[^<>]*\)$", "", cur)   # drop the parameter list, keep template arguments
            counts[cur] = collections.Counter()
            counts[cur]["_insts"] = 0
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            op = m.group(1)
            counts[cur]["_insts"] += 1
            if op in OPS:
                counts[cur][op] += 1
    used = [o for o in OPS if any(c[o] for c in counts.values())]
    print(f"# cuobjdump -sass {os.path.relpath(SO, ROOT)} : {len(counts)} kernels; columns = SASS opcode counts per kernel")
    tot = collections.Counter()
    w = 64
    print(f"{'kernel':{w}s} {'insts':>7s} " + " ".join(f"{o:>7s}" for o in used))
    for k, c in sorted(counts.items()):
        print(f"{k[:w]:{w}s} {c['_insts']:7d} " + " ".join(f"{c[o]:7d}" for o in used))
        tot.update(c)
    print(f"{'TOTAL':{w}s} {tot['_insts']:7d} " + " ".join(f"{tot[o]:7d}" for o in used))


if __name__ == "__main__":
    sys.exit(main())
