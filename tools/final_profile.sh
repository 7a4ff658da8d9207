#!/bin/bash
# GPU box, two calls (results under gpurun_out/ must stay below 64 MiB to travel back):
#   tools/final_profile.sh <tag> tests    the whole GPU test suite, the default bench (without the CPU legs), the ncu launch list
#   tools/final_profile.sh <tag> full <launches per step>   one `ncu --set full` capture per kernel of the last step, summarised ON THE BOX
#                                         (tools/ncu_summary.py -> <tag>_full_kernels.txt + <tag>_traffic_kernels.json), report deleted
# tools/make_profiles.py <tag> then writes profiles/<tag>_{launches.csv,full_summary.txt} + profiles/traffic.json.
tag=${1:-fin}; what=${2:-tests}
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-rank-path"
if [ "$what" = tests ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
  timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv $BENCH > gpurun_out/${tag}_ncu1.log 2>&1
else
  per=${3:?launches per step (from the launch list of the tests call)}
  timeout 600 ncu --set full --clock-control none -s $((per * 4)) -c $per -o /tmp/${tag}_full $BENCH > gpurun_out/${tag}_ncu2.log 2>&1
  python tools/ncu_summary.py /tmp/${tag}_full.ncu-rep gpurun_out/${tag}_full_kernels.txt gpurun_out/${tag}_traffic_kernels.json
fi
ls -la gpurun_out | grep ${tag}
