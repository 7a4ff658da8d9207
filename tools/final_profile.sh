#!/bin/bash
# GPU box: the whole GPU test suite, the default bench (without the CPU legs), then the two ncu passes that
# tools/make_profiles.py turns into profiles/<tag>_{launches.csv,full_summary.txt} + profiles/traffic.json.
# usage (under gpurun): tools/final_profile.sh <tag>
tag=${1:-fin}
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-rank-path"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv $BENCH > gpurun_out/${tag}_ncu1.log 2>&1
# one full capture per kernel of the LAST step: skip the launches of the first four steps
n=$(grep -c '^"[0-9]' gpurun_out/${tag}_launches.csv); per=$((n / 5))
timeout 900 ncu --set full --clock-control none --import-source on -s $((per * 4)) -c $per -o gpurun_out/${tag}_full $BENCH > gpurun_out/${tag}_ncu2.log 2>&1
ls -la gpurun_out | grep ${tag}
