#!/usr/bin/env python
"""Extract per-launch DRAM traffic of the dominant kernel from an `ncu --set full` report into profiles/ncu_traffic.json
(read by bench.py for roofline.traffic).  Usage: ncu_traffic.py report.ncu-rep kernel-substring n_edges out.json"""
import csv, io, json, subprocess, sys
rep, filt, n_edges, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
SC = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    if filt not in d.get("Kernel Name", ""):
        continue
    rd = float(d["dram__bytes_read.sum"].replace(",", "")) * SC[units[hdr.index("dram__bytes_read.sum")]]
    wr = float(d["dram__bytes_write.sum"].replace(",", "")) * SC[units[hdr.index("dram__bytes_write.sum")]]
    json.dump({"kernel": d["Kernel Name"], "n_edges": n_edges, "dram_bytes_read": rd, "dram_bytes_write": wr,
               "traffic": rd + wr, "source": rep.split("/")[-1],
               "gpu_time_ms": float(d["gpu__time_duration.sum"].replace(",", ""))}, open(out, "w"))
    print(open(out).read())
    break
