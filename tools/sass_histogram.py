"""Per-kernel SASS opcode histogram of libmdx_b200.so (cuobjdump -sass): which kernels carry the Blackwell-native
instructions (UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMAPF = TMA, UTCBAR = tcgen05.commit,
SYNCS = mbarrier) and which still run on the legacy tensor path (HMMA = mma.sync).

    python tools/sass_histogram.py > profiles/r02_sass_opcode_histogram.txt
"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "minddiffusion_b200", "libmdx_b200.so")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "LDGSTS", "MUFU",
        "USETMAXREG"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
hist = collections.OrderedDict()
name = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        name = re.sub(r"\(.*", "", name)
        hist[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and name:
        hist[name][m.group(1).split(".")[0]] += 1
        hist[name]["_total"] += 1
print("%-64s %7s " % ("kernel", "instr") + " ".join("%8s" % k for k in KEYS))
agg = collections.Counter()
for name, c in sorted(hist.items(), key=lambda kv: -kv[1]["_total"]):
    if not any(c[k] for k in KEYS[:10]):
        continue
    print("%-64s %7d " % (name[:64], c["_total"]) + " ".join("%8d" % c[k] for k in KEYS))
    agg.update(c)
print("%-64s %7d " % ("TOTAL (kernels listed)", agg["_total"]) + " ".join("%8d" % agg[k] for k in KEYS))
print("\nkernels with HMMA (mma.sync legacy tensor path): " +
      ", ".join(sorted(n for n, c in hist.items() if c["HMMA"])) or "none")
