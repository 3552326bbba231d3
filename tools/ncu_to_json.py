#!/usr/bin/env python
"""Extract the per-launch counters bench.py quotes from an .ncu-rep of the potential kernel.
usage: python tools/ncu_to_json.py gpurun_out/prof_pot.ncu-rep 1000000 > profiles/potential_ncu.json"""
import csv, json, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
def val(k):
    i = hdr.index(k)
    v = float(r[i].replace(",", ""))
    u = units[i]
    return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9}.get(u, 1)
print(json.dumps({
    "kernel": r[hdr.index("Kernel Name")], "stars": int(sys.argv[2]),
    "dram_read_bytes": val("dram__bytes_read.sum"), "dram_write_bytes": val("dram__bytes_write.sum"),
    "duration_s_under_ncu": val("gpu__time_duration.sum"),
    "tensor_active_pct": float(r[hdr.index("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed")]),
    "issue_active_pct": float(r[hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")]),
    "registers": int(float(r[hdr.index("launch__registers_per_thread")])),
}, indent=1))
