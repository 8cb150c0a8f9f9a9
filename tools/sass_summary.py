#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in the shipped library (B200_PROFILING.md: UTC*MMA =
tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA, HMMA = legacy mma.sync).
usage: python tools/sass_summary.py [lib.so] > profiles/r02_sass_summary.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "lipvq_vae_b200/_lib/liblipvq_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "SYNCS", "REDG", "ATOMG", "MUFU"]
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for n in names:
            if op.startswith(n):
                counts[cur][n] += 1
print("# SASS mnemonic counts per kernel, %s (cuobjdump -sass; sm_100a)" % lib)
print("%-72s %7s " % ("kernel", "instrs") + " ".join("%8s" % n for n in names))
tot = collections.Counter()
for k, c in counts.items():
    print("%-72s %7d " % (k[:72], c["_total"]) + " ".join("%8d" % c[n] for n in names))
    tot.update(c)
print("%-72s %7d " % ("TOTAL", tot["_total"]) + " ".join("%8d" % tot[n] for n in names))
