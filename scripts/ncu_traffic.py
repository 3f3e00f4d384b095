"""DRAM traffic per launch of one kernel from an `ncu --set full` report -> the JSON bench.py reads for roofline.traffic.
  python scripts/ncu_traffic.py report.ncu-rep <workload> <entry point> <kernel substring> "<source note>" > profiles/rN_traffic.json"""
import csv
import io
import json
import subprocess
import sys

rep, workload, entry, pat, note = sys.argv[1:6]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out)))
h, units = r[0], r[1]
unit = dict(zip(h, units))
SC = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
rd = wr = us = 0.0
n = 0
for x in r[2:]:
    d = dict(zip(h, x))
    if pat not in d["Kernel Name"]:
        continue
    n += 1
    rd += float(d["dram__bytes_read.sum"]) * SC[unit["dram__bytes_read.sum"]]
    wr += float(d["dram__bytes_write.sum"]) * SC[unit["dram__bytes_write.sum"]]
    us += float(d["gpu__time_duration.sum"]) * SC[unit["gpu__time_duration.sum"]]
print(json.dumps({"source": note, workload: {entry: {"launches": n, "dram_read_MB": round(rd / 1e6, 1), "dram_write_MB": round(wr / 1e6, 1),
                                                    "traffic_bytes_per_launch": int((rd + wr) / max(n, 1)),
                                                    "time_us_under_ncu": round(us, 1)}}}, indent=1))
