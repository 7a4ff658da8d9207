#!/usr/bin/env python
"""Per-CUDA-source-line instruction / stall-sample shares from an .ncu-rep (needs -lineinfo + --import-source on).
usage: tools/ncu_lines.py report.ncu-rep [kernel-index] [min-share-percent]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.6
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "File Path":
        cur = []; blocks.append(cur); continue
    if cur is not None: cur.append(r)
blk = blocks[which]
print(blk[0])
hdr = blk[1]
ii, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
lines = [r for r in blk[2:] if len(r) > ii and r[0].isdigit() and r[2] == "-"]
tot = sum(int(r[ii]) for r in lines); ts = sum(int(r[si]) for r in lines)
print("total warp-inst", tot, "samples", ts)
for r in lines:
    if int(r[ii]) > tot * thr / 100 or int(r[si]) > ts * thr / 100:
        print(f"L{r[0]:>4s} inst={int(r[ii])/tot*100:5.1f}% thr={int(r[ti])/max(int(r[ii]),1):5.1f} samp={int(r[si])/ts*100:5.1f}% | {r[1].strip()[:105]}")
