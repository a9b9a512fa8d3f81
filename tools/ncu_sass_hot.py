#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv`: per kernel, warp-instructions executed by opcode and the hottest
SASS instructions with their stall samples.  Usage: ncu_sass_hot.py report.ncu-rep [kernel-substring] [top-n]"""
import csv, subprocess, sys, collections, io
rep = sys.argv[1]; filt = sys.argv[2] if len(sys.argv) > 2 else ""; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
kern = None; hdr = None; data = collections.OrderedDict()
for r in rows:
    if len(r) >= 2 and r[0] == "Kernel Name":
        kern = r[1]; data[kern] = []; hdr = None; continue
    if r and r[0] == "Address":
        hdr = r; continue
    if kern and hdr and len(r) == len(hdr):
        data[kern].append(dict(zip(hdr, r)))
for k, ins in data.items():
    if filt not in k: continue
    tot = sum(int(i["Instructions Executed"]) for i in ins)
    samp = sum(int(i["# Samples"]) for i in ins)
    print(f"== {k}: {tot} warp-instr, {samp} samples, {len(ins)} SASS lines")
    by = collections.Counter(); bys = collections.Counter()
    for i in ins:
        op = i["Source"].split()[0] if not i["Source"].strip().startswith("@") else i["Source"].split()[1]
        op = op.split(".")[0]
        by[op] += int(i["Instructions Executed"]); bys[op] += int(i["# Samples"])
    for op, c in by.most_common(25):
        print(f"  {op:10s} {c:12d} {100*c/tot:5.1f}%  samples {100*bys[op]/max(samp,1):5.1f}%")
    print("  -- hottest by samples")
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for i in sorted(ins, key=lambda i: -int(i["# Samples"]))[:topn]:
        st = sorted(((int(i[c]), c) for c in stall_cols), reverse=True)[:2]
        print(f"  {int(i['# Samples']):7d} {int(i['Instructions Executed']):10d}  {i['Source'].strip()[:70]:70s} {st}")
