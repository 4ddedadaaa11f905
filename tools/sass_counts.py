"""Per-kernel counts of the SASS mnemonics that show the kernels are written for sm_100a (packed FP32, TMA, mbarrier, texture),
from `cuobjdump -sass` of the built library.  usage: python tools/sass_counts.py [libvolrt.so] > profiles/rNN_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "volumetricrenderer_b200", "libvolrt.so")
WATCH = ["FFMA2", "FMUL2", "FADD2", "FFMA", "IMAD", "LOP3", "PRMT", "LDS", "LDG", "TEX", "TLD", "UTMALDG", "SYNCS", "MUFU", "FRND", "F2I",
         "ATOMG", "RED", "REDUX", "VOTE", "SHFL", "BAR", "LDL", "STL"]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kernels, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(volrt::MarchParams\)|volrt::|\(anonymous namespace\)::", "", cur)
        kernels[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        kernels[cur]["_total"] += 1
        kernels[cur][m.group(1)] += 1
print("# static SASS instruction counts per kernel, cuobjdump -sass %s (sm_100a)" % os.path.relpath(lib, ROOT))
print("# FFMA2/FMUL2/FADD2 = packed FP32 (two binary32 operations per instruction, sm_100+), UTMALDG = TMA tensor load, SYNCS = mbarrier,")
print("# TEX/TLD = texture unit, PRMT = byte permute, REDUX = warp reduce")
hdr = ["total"] + WATCH
print("%-78s %s" % ("kernel", " ".join("%7s" % h for h in hdr)))
for k, c in kernels.items():
    vals = [c["_total"]] + [c[w] for w in WATCH]
    print("%-78s %s" % (k[:78], " ".join("%7d" % v for v in vals)))
tot = collections.Counter()
for c in kernels.values():
    tot.update(c)
print("%-78s %s" % ("ALL KERNELS", " ".join("%7d" % v for v in [tot["_total"]] + [tot[w] for w in WATCH])))
