#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in the built library (no GPU needed):
   python tools/sass_summary.py > profiles/sass_summary_rN.txt
tcgen05.mma = UTCHMMA(.2CTA), tcgen05.ld/st = LDTM/STTM, TMA tensor copies = UTMALDG/UTMASTG/UTMAREDG, 1-D bulk TMA = UBLKCP,
mbarrier = SYNCS; HMMA (legacy mma.sync) must be absent."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "avion_b200", "csrc", "libavion_b200.so")
MN = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "HMMA", "MUFU.EX2", "F2FP",
      "FFMA2", "FADD2", "FMUL2", "REDG", "RED", "ATOMG"]

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
names = {}
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        raw = m.group(1)
        if raw not in names:
            d = subprocess.run(["c++filt", raw], capture_output=True, text=True).stdout.strip()
            d = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", d)
            names[raw] = re.sub(r"\(.*", "", d).replace("void ", "")
        cur = names[raw]
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m:
        continue
    op = m.group(1)
    for k in MN:
        if op == k or op.startswith(k + "."):
            if k == "UTCHMMA" and op.startswith("UTCHMMA.2CTA"):
                continue
            if k == "RED" and op.startswith("REDG"):
                continue
            counts[cur][k] += 1
print("SASS summary of", os.path.relpath(LIB, ROOT), "(cuobjdump -sass, sm_100a), static instruction counts per kernel")
print(__doc__.split("\n", 2)[2].strip())
print()
tot = collections.Counter()
for k, c in counts.items():
    if not c:
        continue
    tot.update(c)
    print(f"{k[:78]:78s} " + " ".join(f"{m}={c[m]}" for m in MN if c[m]))
print()
print("TOTAL " + " ".join(f"{m}={tot[m]}" for m in MN))
