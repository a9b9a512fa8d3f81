#!/usr/bin/env python
"""profiles/ncu_traffic.json from a tools/ncu_summary.py text (the report itself stays on the GPU box): DRAM bytes of
the first captured launch.  usage: ncu_traffic_txt.py <tag>_fill_full.txt [n_edges] > ncu_traffic.json"""
import json, re, sys
SC = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
T = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
kern, vals = None, {}
for line in open(sys.argv[1]):
    if line.startswith("kernel:"):
        if kern: break
        kern = line.split("kernel:", 1)[1].strip()
        continue
    m = re.match(r"\s+(\S+)\s+([0-9.,]+)\s+(\S+)", line)
    if m: vals[m.group(1)] = (float(m.group(2).replace(",", "")), m.group(3))
rd = vals["dram__bytes_read.sum"][0] * SC[vals["dram__bytes_read.sum"][1]]
wr = vals["dram__bytes_write.sum"][0] * SC[vals["dram__bytes_write.sum"][1]]
t = vals["gpu__time_duration.sum"]
print(json.dumps({"kernel": kern, "n_edges": int(sys.argv[2]) if len(sys.argv) > 2 else 254388598, "dram_bytes_read": rd,
                  "dram_bytes_write": wr, "traffic": rd + wr, "source": sys.argv[1].split("/")[-1],
                  "gpu_time_ms": t[0] * T.get(t[1], 1.0)}))
