#!/usr/bin/env python
"""profiles/<tag>_launches.csv + <tag>_full_summary.txt + traffic.json from what the two calls of tools/final_profile.sh
left in gpurun_out/: <tag>_launches.csv (ncu launch list), <tag>_full_kernels.txt + <tag>_traffic_kernels.json (the
`ncu --set full` capture summarised on the GPU box by tools/ncu_summary.py; the report itself is too large to travel).
usage: tools/make_profiles.py <tag>"""
import csv, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
go, pr = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
shutil.copy(os.path.join(go, f"{tag}_launches.csv"), os.path.join(pr, f"{tag}_launches.csv"))
rows = [r for r in csv.reader(open(os.path.join(go, f"{tag}_launches.csv"))) if len(r) > 10 and r[0].isdigit()]
steps = sum(1 for r in rows if "scan_kernel" in r[4])  # one scan pass per step (config 2): warm-up, stage-timed and plain loops
per = len(rows) // steps
last = rows[-per:]
tot = sum(int(r[-1]) for r in last)
out = os.path.join(pr, f"{tag}_full_summary.txt")
with open(out, "w") as f:
    f.write(f"# {tag}: python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-rank-path  (config 2: 200 MiB text, 1000 queries; {steps} steps in the run)\n")
    f.write(f"# launch list of the last step ({per} launches, {tot/1e6:.3f} ms of kernel time; ncu --metrics gpu__time_duration.sum --clock-control none)\n")
    for r in last:
        f.write(f"#  {int(r[-1])/1e3:9.1f} us {int(r[-1])/tot*100:5.1f}%  {r[4][:70]} grid={r[8]} block={r[7]}\n")
    f.write("# ncu --set full --clock-control none, the kernels of the last steps of the same command (a kernel may appear twice):\n")
    f.write(open(os.path.join(go, f"{tag}_full_kernels.txt")).read())
k = json.load(open(os.path.join(go, f"{tag}_traffic_kernels.json")))
def g(name):
    for kk, v in k.items():
        if name in kk: return sum(v) / len(v)
    return None
stage = {"scan_count": g("scan_kernel"), "scan_fill": g("scan_scatter"), "join_probe": (g("join_probe") or 0) + (g("join_rev_count") or 0),
         "join_compact": (g("join_compact") or 0) + (g("join_rev_expand") or 0), "select_tile": (g("sel_succ") or 0) + (g("sel_tile") or 0),
         "select_walk": g("sel_mark"), "emit": g("sel_emit")}
sys.path.insert(0, ROOT)
import bench  # noqa: E402
json.dump({"source": f"profiles/{tag}_full_summary.txt (dram__bytes_read.sum + dram__bytes_write.sum per launch)",
           "source_hash": bench.kernel_source_hash(), **stage},
          open(os.path.join(pr, "traffic.json"), "w"), indent=1)
print(open(out).read()[:3000])
