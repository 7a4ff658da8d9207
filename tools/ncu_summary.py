#!/usr/bin/env python
"""Per-kernel summary of an .ncu-rep (--set full): duration, DRAM traffic, throughput, issue, occupancy.
usage: tools/ncu_summary.py report.ncu-rep [out.txt] [traffic.json]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__waves_per_multiprocessor", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "launch__grid_size", "launch__block_size"]
idx = {h: i for i, h in enumerate(hdr)}
def tobytes(v, u):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
out, traffic = [], {}
for r in data:
    name = r[idx["Kernel Name"]]
    out.append(f"Kernel Name  {name}")
    for w in want:
        if w in idx:
            out.append(f"  {w:66s} {r[idx[w]]} {units[idx[w]]}")
    if "dram__bytes_read.sum" in idx:
        t = tobytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) + tobytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
        traffic.setdefault(name.split("(")[0].replace("void ", ""), []).append(t)
    out.append("---")
txt = "\n".join(out)
if len(sys.argv) > 2: open(sys.argv[2], "a").write(txt + "\n")
else: print(txt)
if len(sys.argv) > 3: json.dump(traffic, open(sys.argv[3], "w"), indent=1)
