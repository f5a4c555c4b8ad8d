"""Per-launch DRAM traffic and utilisation of the fused histogram kernel from an `ncu --set full` report
-> profiles/hist_kernel_traffic.json (read by bench.py for roofline.traffic).

usage: python tools/ncu_traffic.py gpurun_out/prof_hist.ncu-rep > profiles/hist_kernel_traffic.json
"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    u = dict(zip(hdr, units))
    best = None
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        if "hist_kernel" in d.get("Kernel Name", ""):
            best = d
    d = best

    def val(k):
        return float(d[k]) * UNIT.get(u[k], 1)

    # the capture is only meaningful for the kernel source it was taken from: stamp it with `git hash-object hist.cu`
    src = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "opencv_b200", "csrc", "hist.cu")
    data = open(src, "rb").read()
    out = {"source": rep, "kernel": d["Kernel Name"][:80],
           "source_hash": {"opencv_b200/csrc/hist.cu": hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()},
           "dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
           "dram_read_bytes": val("dram__bytes_read.sum"), "dram_write_bytes": val("dram__bytes_write.sum"),
           "ncu": {"duration_us_under_ncu": float(d["gpu__time_duration.sum"]),
                   "issue_active_pct": float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]),
                   "fma_pipe_pct": float(d["sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"]),
                   "alu_pipe_pct": float(d["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]),
                   "lsu_pipe_pct": float(d["sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]),
                   "shared_wavefronts": float(d["l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]),
                   "warp_instructions": float(d["smsp__inst_executed.sum"]),
                   "dram_throughput_pct": float(d["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"])}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
