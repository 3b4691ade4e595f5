#!/bin/bash
# Run ON the GPU box (via gpurun): plain bench first, then the ncu launch list and one full capture of the top kernel.
# Usage: tools/gpu_profile.sh <tag> [kernel-regex]
set -u
TAG=${1:-r01}
KRE=${2:-conv3x3_igemm}
CMD="python bench.py --steps 2 --warmup 3 --no-extras"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s 700 -c 3 \
    -o gpurun_out/${TAG}_${KRE}_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo done
