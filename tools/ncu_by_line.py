"""Per-source-line instruction budget of one kernel: joins the SASS page of an .ncu-rep with nvdisasm's line info
of the object the library was linked from (same build), so every executed warp-instruction, stall sample and
shared-memory wavefront is attributed to a line of the .cu file.
usage: python tools/ncu_by_line.py X.ncu-rep build/hist.o <kernel-substring> [units_per_launch] [--ranges a-b:name,...]"""
import csv, io, os, re, subprocess, sys, tempfile
from collections import defaultdict

rep, obj, kern = sys.argv[1:4]
units = float(sys.argv[4]) if len(sys.argv) > 4 and not sys.argv[4].startswith("--") else None
ranges = []
for a in sys.argv[4:]:
    if a.startswith("--ranges="):
        for part in a[9:].split(","):
            r, name = part.split(":"); lo, hi = r.split("-"); ranges.append((int(lo), int(hi), name))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
line_of = {}
cur = None; active = False
for l in sass:
    if l.startswith(".text.") and l.rstrip().endswith(":"):
        active = kern in l
    if not active:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(txt) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(txt[start:]))))
base = int(rows[0]["Address"], 16)
agg = defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows:
    off = int(r["Address"], 16) - base
    key = line_of.get(off, ("?", 0))
    v = (int(r["Instructions Executed"]), int(r["# Samples"]), int(r["L1 Wavefronts Shared"]))
    for i in range(3):
        agg[key][i] += v[i]; tot[i] += v[i]
print("total: warp-inst %d samples %d smem-wavefronts %d" % tuple(tot))
if ranges:
    rg = defaultdict(lambda: [0, 0, 0])
    for (f, ln), v in agg.items():
        name = next((n for lo, hi, n in ranges if lo <= ln <= hi), "other")
        for i in range(3):
            rg[name][i] += v[i]
    print("%-28s %12s %7s %8s %7s %12s %7s %s" % ("section", "warp-inst", "%", "samples", "%", "wavefronts", "%", "thread-inst/unit"))
    for n, v in sorted(rg.items(), key=lambda kv: -kv[1][0]):
        print("%-28s %12d %6.2f%% %8d %6.2f%% %12d %6.2f%% %s" % (n, v[0], 100. * v[0] / tot[0], v[1], 100. * v[1] / max(tot[1], 1), v[2],
              100. * v[2] / max(tot[2], 1), "%.2f" % (32 * v[0] / units) if units else ""))
else:
    for (f, ln), v in sorted(agg.items(), key=lambda kv: kv[0][1]):
        print("%s:%-5d %12d %6.2f%% %8d %6.2f%% %12d" % (f, ln, v[0], 100. * v[0] / tot[0], v[1], 100. * v[1] / max(tot[1], 1), v[2]))
