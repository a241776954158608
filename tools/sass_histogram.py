"""Per-kernel histogram of the SASS opcodes that show what the kernels are made of (cuobjdump -sass of the built library):
UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UBLKCP (cp.async.bulk 1-D), UTMALDG (tensor-map TMA), UTCBAR
(tcgen05.commit), UTCATOMSWS (TMEM alloc), SYNCS (mbarrier), REDG (red.global) / ATOM, HMMA / IMMA (legacy warp MMA -- expected absent).
  python tools/sass_histogram.py > profiles/sass_opcodes_r2.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "deep-go_b200", "libdeepgo_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WATCH = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTCBAR", "UTCATOMSWS", "SYNCS", "REDG", "ATOM", "HMMA", "IMMA", "FFMA", "LDG", "STG", "LDS", "STS"]
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        op = m.group(1)
        hist[kern]["_all"] += 1
        for w in WATCH:
            if op == w or op.startswith(w + "."):
                hist[kern][w] += 1
print("SASS opcode histogram of %s (sm_100a), built from commit %s" % (os.path.basename(lib), subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()))
print("%-60s %8s " % ("kernel", "instrs") + " ".join("%7s" % w[:7] for w in WATCH))
tot = collections.Counter()
for k, c in hist.items():
    print("%-60s %8d " % (k[:60], c["_all"]) + " ".join("%7d" % c[w] for w in WATCH))
    tot.update(c)
print("%-60s %8d " % ("TOTAL", tot["_all"]) + " ".join("%7d" % tot[w] for w in WATCH))
