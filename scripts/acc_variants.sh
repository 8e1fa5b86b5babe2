#!/bin/bash
# DRAM traffic and time of the bucket-accumulation kernel at MSM 2^22 under different L2 fetch-granularity hints
# (H2B_L2_FETCH -> cudaLimitMaxL2FetchGranularity): plain timing first, then one ncu pass with the DRAM byte counters.
mkdir -p gpurun_out
for v in 32 64 128; do
  H2B_L2_FETCH=$v python scripts/sweep.py --msm 22,24 --ntt "" --poly "" > gpurun_out/l2f_$v.json 2> gpurun_out/l2f_$v.err || tail -3 gpurun_out/l2f_$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/l2f_$v.json"))
for k,r in d["msm"].items(): print("l2fetch $v k",k,"device_ms",round(r["device_ms"],3),"acc",round(r["kernels_ms"]["msm_accumulate"],3))
PY
  H2B_L2_FETCH=$v ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:msm_accumulate_kernel -c 1 --csv --log-file gpurun_out/l2f_ncu_$v.csv python scripts/prof_target.py 22 0 > /dev/null 2>&1
  grep -E "dram__bytes|gpu__time" gpurun_out/l2f_ncu_$v.csv | awk -F, '{print $(NF-2), $(NF-1), $NF}' | head -3
done
