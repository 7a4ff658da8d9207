#!/usr/bin/env python
"""Per-CUDA-source-line instruction / stall-sample shares from an .ncu-rep (needs -lineinfo + --import-source on).
usage: tools/ncu_lines.py report.ncu-rep [kernel-substring] [min-share-percent]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; kern = sys.argv[2] if len(sys.argv) > 2 else ""; thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.6
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "File Path":
        cur = {"file": r[1], "rows": []}; blocks.append(cur); continue
    if cur is not None:
        if r and r[0] == "Function Name": cur["fn"] = r[1]; continue
        cur["rows"].append(r)
fns = []
for b in blocks:
    if b.get("fn") not in fns: fns.append(b.get("fn"))
fn = next((f for f in fns if kern in f), fns[0])
sel = [b for b in blocks if b.get("fn") == fn]
print(fn)
allrows = []
for b in sel:
    hdr = b["rows"][0]
    ii, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    for r in b["rows"][1:]:
        if len(r) > ii and r[0].isdigit() and r[2] == "-":
            allrows.append((b["file"].split("/")[-1], int(r[0]), r[1], int(r[ii]), int(r[ti]), int(r[si])))
tot = sum(r[3] for r in allrows); ts = sum(r[5] for r in allrows)
print("total warp-inst", tot, "samples", ts)
for f, ln, src, ins, tins, smp in allrows:
    if ins > tot * thr / 100 or smp > ts * thr / 100:
        print(f"{f[:10]:10s} L{ln:>4d} inst={ins/tot*100:5.1f}% thr={tins/max(ins,1):5.1f} samp={smp/ts*100:5.1f}% | {src.strip()[:100]}")
