"""Condense `ncu -i X.ncu-rep --page source --csv` (SASS view) into a per-opcode table: executed warp-instructions,
stall samples and shared-memory wavefronts per opcode class -- the instruction budget of a kernel.
usage: python tools/ncu_source_digest.py X.ncu-rep [pixels_or_units_per_launch]"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
units = float(sys.argv[2]) if len(sys.argv) > 2 else None
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = txt.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
print(lines[start - 1][:200] if start else "")
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
tot_inst = sum(int(r["Instructions Executed"]) for r in rows)
tot_thr = sum(int(r["Thread Instructions Executed"]) for r in rows)
tot_samp = sum(int(r["# Samples"]) for r in rows)
tot_wf = sum(int(r["L1 Wavefronts Shared"]) for r in rows)
by = defaultdict(lambda: [0, 0, 0, 0, 0])
for r in rows:
    src = r["Source"].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2) if m else src.split()[0]
    key = op.split(".")[0]
    if key in ("LDS", "STS", "LDG", "STG", "MUFU", "SHFL", "ATOMG", "LDC"):
        key = ".".join(op.split(".")[:3]) if key in ("LDS", "STS", "LDG", "STG") else op
    b = by[key]
    b[0] += int(r["Instructions Executed"]); b[1] += int(r["# Samples"]); b[2] += int(r["L1 Wavefronts Shared"])
    b[3] += int(r["L1 Wavefronts Shared Ideal"]); b[4] += 1
print("total warp-instructions %d  thread-instructions %d  samples %d  shared wavefronts %d  SASS lines %d" % (tot_inst, tot_thr, tot_samp, tot_wf, len(rows)))
if units:
    print("per unit: %.2f thread-instructions, %.3f shared wavefronts x32" % (tot_thr / units, 32 * tot_wf / units))
print("%-22s %12s %7s %8s %7s %12s %12s %5s" % ("opcode", "warp-inst", "%inst", "samples", "%samp", "smem wavefr", "ideal", "lines"))
for k, b in sorted(by.items(), key=lambda kv: -kv[1][0]):
    if b[0] == 0 and b[1] == 0:
        continue
    print("%-22s %12d %6.2f%% %8d %6.2f%% %12d %12d %5d" % (k, b[0], 100.0 * b[0] / tot_inst, b[1], 100.0 * b[1] / max(tot_samp, 1), b[2], b[3], b[4]))
