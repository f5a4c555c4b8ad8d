"""SASS digest of the shipped library: per kernel, the instruction mnemonics that show which Blackwell features it uses
(FFMA2 = packed FP32, UBLKCP = bulk-copy engine, SYNCS = mbarrier, IDP = integer dot product, UTC*MMA/LDTM = tcgen05
-- none expected on this path) plus registers / shared memory / spills from the resource usage dump.
usage: python tools/sass_digest.py [opencv_b200/libhogb200.so] > profiles/rNN_sass_digest.txt"""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "opencv_b200", "libhogb200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "--dump-resource-usage", so], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
usage = {}
cur = None
for l in res.splitlines():
    m = re.search(r"Function (\S+):", l)
    if m:
        cur = m.group(1); continue
    if cur and "REG:" in l:
        usage[cur] = l.strip(); cur = None
kern = OrderedDict(); cur = None
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1); kern[cur] = Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)(\.[A-Z0-9_.]+)?", l)
    if m and cur:
        op = m.group(1); kern[cur][op] += 1; kern[cur]["_total"] += 1
        if op in ("LDS", "STS", "LDG", "STG", "LDL", "STL"):
            kern[cur][op + (m.group(2) or "")] += 1
KEYS = ["_total", "FFMA2", "FMUL2", "FADD2", "FFMA", "IDP", "UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "LDL", "STL", "MUFU", "SHFL",
        "ATOMG", "ATOMS", "REDG", "HMMA", "UTCHMMA", "UTCQMMA", "LDTM", "UTMALDG"]
print("# %s  (cuobjdump -sass, static instruction counts per kernel; arch sm_100a)" % os.path.relpath(so))
print("%-64s %s" % ("kernel", " ".join("%7s" % k.strip("_") for k in KEYS)))
for k, c in kern.items():
    name = demangle(k)
    if name.endswith(")"):                      # drop the parameter list: the '(' that matches the final ')'
        depth = 0
        for i in range(len(name) - 1, -1, -1):
            depth += name[i] == ")"; depth -= name[i] == "("
            if depth == 0:
                name = name[:i]; break
    name = name.replace("hogk::(anonymous namespace)::", "").replace("hogk::<unnamed>::", "").replace("void ", "").replace("(int)", "").replace("(bool)", "")
    print("%-64s %s" % (name[:64], " ".join("%7d" % c.get(x, 0) for x in KEYS)))
    if k in usage:
        print("    %s" % usage[k])
    wide = {x: n for x, n in c.items() if "." in x and x.split(".")[0] in ("LDS", "STS", "LDG", "STG")}
    if wide:
        print("    " + "  ".join("%s=%d" % kv for kv in sorted(wide.items())))
