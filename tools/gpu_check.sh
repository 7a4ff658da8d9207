#!/bin/bash
# GPU-box check: parity tests, then the bench stage table.  usage: tools/gpu_check.sh [tag]
tag=${1:-x}
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$tag.json"))
print("step ms", round(d["ms_per_step"],3), "e2e ms", round(d["e2e"]["ms_per_step"],3), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"],4))
for k,v in d["stages"].items(): print(f"  {k:14s} {v['ms_per_step']:.3f} ms  {v['GBps']:.0f} GB/s")
PY
