#!/usr/bin/env python
"""SASS op counts per kernel of the built library: the mnemonics that prove what the kernels are made of
(TMA: UTMASTG / UBLKCP; cp.async: LDGSTS; mbarrier: SYNCS; packed fp32 adds: FADD2; NO FFMA / FFMA2 in the pixel stage --
the no-contraction exactness recipe; FRND / F2IP / I2F: the XU-pipe share).
    python scripts/dev/sass_summary.py [lib.so] > profiles/r02_sass_ops.txt"""
import collections
import re
import subprocess
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "jpeg_decoder_b200", "libb200jpeg.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
OPS = ["UTMASTG", "UBLKCP", "LDGSTS", "SYNCS", "FADD2", "FADD", "FMUL", "FFMA", "FFMA2", "FRND", "F2IP", "I2F", "LDS", "STS", "LDG", "STG",
       "SHFL", "VOTE", "BAR", "ATOM", "RED", "IMAD", "LOP3", "SHF", "ISETP", "SEL"]
print("cubins:", ", ".join(arch))
print("%-28s %6s " % ("kernel", "total") + " ".join("%7s" % o for o in OPS))
cur, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0].replace("b2::", "").replace("void ", "")
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        counts[cur]["total"] += 1
        counts[cur][m.group(1)] += 1
for k, c in counts.items():
    print("%-28s %6d " % (k[:28], c["total"]) + " ".join("%7d" % c[o] for o in OPS))
