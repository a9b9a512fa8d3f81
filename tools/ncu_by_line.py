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
# per-function summary: every source line is attributed to the nearest preceding function definition in its file
import os, re, bisect
CS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "defect-gnn-cpp_b200", "csrc")
DEF = re.compile(r"^(?:template.*>\s*)?(?:static\s+)?(?:__global__|__device__|__host__|DG_HD|inline).*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(")
fmap = {}
for fn_ in os.listdir(CS):
    if not fn_.endswith((".cu", ".cuh", ".h")):
        continue
    starts, names = [], []
    for ln, text in enumerate(open(os.path.join(CS, fn_), errors="ignore"), 1):
        m = DEF.match(text)
        if m and not text.rstrip().endswith(";"):
            starts.append(ln); names.append(m.group(1))
    fmap[fn_] = (starts, names)
ph = collections.OrderedDict()
for (f, l), a in agg.items():
    name = "other:" + f
    if f in fmap and fmap[f][0]:
        k = bisect.bisect_right(fmap[f][0], l) - 1
        if k >= 0:
            name = f"{f}:{fmap[f][1][k]}"
    v = ph.setdefault(name, [0, 0]); v[0] += a[0]; v[1] += a[1]
print("-- by function (source line -> nearest preceding definition)")
for n, v in sorted(ph.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{n:48s} instr {100*v[0]/ti:5.1f}%  samples {100*v[1]/ts:5.1f}%")
