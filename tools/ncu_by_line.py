#!/usr/bin/env python
"""Aggregate `ncu --page source --print-source sass,cuda` by CUDA source line: warp instructions and stall samples.
Usage: ncu_by_line.py report.ncu-rep <kernel-substring> [top-n]"""
import csv, io, subprocess, sys, collections
rep, filt = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 50
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fn = None; fpath = None; hdr = None
agg = collections.defaultdict(lambda: [0, 0, ""])
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fpath = r[1]; continue
    if len(r) >= 2 and r[0] == "Function Name": fn = r[1]; hdr = None; continue
    if r and r[0] == "Line No": hdr = r; continue
    if not (fn and hdr and filt in fn) or len(r) != len(hdr): continue
    if r[0] == "": continue  # SASS sub-row
    ie = hdr.index("Instructions Executed"); sa = hdr.index("# Samples")
    key = (fpath.split("/")[-1], int(r[0]))
    a = agg[key]; a[0] += int(r[ie]); a[1] += int(r[sa]); a[2] = r[1].strip()[:90]
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print(f"{filt}: {ti} warp-instr, {ts} samples")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f"{k[0]:16s}:{k[1]:4d} instr {100*a[0]/ti:5.1f}% samp {100*a[1]/ts:5.1f}%  {a[2]}")
# phase summary for the fill kernel (line ranges of row_sort.cuh / rbf_math.cuh / dgnn_direct.cu)
PH = [("row_sort.cuh", 73, 112, "rbf epilogue"), ("row_sort.cuh", 113, 150, "bitonic: sqrt+key"),
      ("row_sort.cuh", 151, 195, "bitonic: network"), ("row_sort.cuh", 196, 205, "bitonic: clash check"),
      ("row_sort.cuh", 206, 222, "bitonic: emit"), ("row_sort.cuh", 223, 250, "sort dispatch/init"),
      ("row_sort.cuh", 251, 999, "counting sort (fallback)"),
      ("rbf_math.cuh", 0, 999, "rbf epilogue"), ("dgnn_direct.cu", 0, 9999, "traversal"), ("geom.h", 0, 9999, "traversal")]
ph = collections.OrderedDict()
for (f, l), a in agg.items():
    name = next((p[3] for p in PH if p[0] == f and p[1] <= l <= p[2]), "other:" + f)
    v = ph.setdefault(name, [0, 0]); v[0] += a[0]; v[1] += a[1]
print("-- phases")
for n, v in sorted(ph.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:28s} instr {100*v[0]/ti:5.1f}%  samples {100*v[1]/ts:5.1f}%")
