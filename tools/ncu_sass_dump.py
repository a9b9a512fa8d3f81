#!/usr/bin/env python
"""Dump `ncu --page source` SASS of one kernel in address order with exec counts: ncu_sass_dump.py rep kernel-substr [min_exec]"""
import csv, subprocess, sys, io
rep, filt = sys.argv[1], sys.argv[2]; mn = int(sys.argv[3]) if len(sys.argv) > 3 else 0
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
kern = None; hdr = None
for r in csv.reader(io.StringIO(txt)):
    if len(r) >= 2 and r[0] == "Kernel Name": kern = r[1]; hdr = None; continue
    if r and r[0] == "Address": hdr = r; continue
    if kern and filt in kern and hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if int(d["Instructions Executed"]) >= mn:
            print(f'{d["Address"][-5:]} {int(d["Instructions Executed"]):10d} {int(d["# Samples"]):6d}  {d["Source"].strip()}')
