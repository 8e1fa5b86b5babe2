#!/bin/bash
# usage: scripts/ncu_capture.sh <tag> <kernel-regex> <count> <msm_k> <ntt_k>
# Plain run first (must exit 0), then one `ncu --set full` capture; the report is exported to CSV pages on the box
# and dropped if it is too large to travel back (gpurun_out/ is capped at 64 MiB: NCU_KEEP_BYTES, NCU_SOURCE=0 skips the source page).
set -u
tag=$1; regex=$2; count=$3; mk=$4; nk=$5
mkdir -p gpurun_out
python scripts/prof_target.py $mk $nk > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain_$tag.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$regex" -c $count -o gpurun_out/$tag python scripts/prof_target.py $mk $nk > gpurun_out/ncu_$tag.log 2>&1
ncu -i gpurun_out/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/$tag.ncu-rep --page details --csv > gpurun_out/${tag}_details.csv 2>/dev/null
if [ "${NCU_SOURCE:-1}" = "1" ]; then ncu -i gpurun_out/$tag.ncu-rep --page source --csv > gpurun_out/${tag}_source.csv 2>/dev/null; fi
ls -la gpurun_out/
sz=$(stat -c %s gpurun_out/$tag.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -gt ${NCU_KEEP_BYTES:-30000000} ]; then rm -f gpurun_out/$tag.ncu-rep; echo "report too large ($sz), kept CSV pages only"; fi
tail -n 3 gpurun_out/plain_$tag.log gpurun_out/ncu_$tag.log
