#!/bin/bash
# usage (on the GPU box, from the repo root): scripts/capture_traffic.sh <commit> [msm_k] [ntt_k]
# DRAM bytes and duration of EVERY launch of one MSM and one NTT at the headline size, for bench.py's `roofline.traffic`.
# The plain run must exit 0 first; numbers printed under ncu are never bench values.
set -u
commit=${1:-unknown}; mk=${2:-24}; nk=${3:-24}
mkdir -p gpurun_out
H2B_PROF_REPS=1 python scripts/prof_target.py $mk $nk > gpurun_out/traffic_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/traffic_plain.log; exit 1; }
H2B_PROF_REPS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/traffic.csv python scripts/prof_target.py $mk $nk > gpurun_out/traffic_ncu.log 2>&1
python scripts/traffic_to_json.py gpurun_out/traffic.csv gpurun_out/full_size_dram_traffic.json "$commit"
