#!/bin/bash
# Run ON the GPU box (via gpurun): GPU tests, plain bench, kernel breakdown, ncu launch list (eager launches, so that
# every kernel is a separate ncu result) and one --set full capture of the top kernel.
# Usage: tools/gpu_session.sh <tag> [kernel-regex]
set -u
TAG=${1:-r01}
KRE=${2:-conv3x3_igemm}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || { echo "bench failed"; tail -20 gpurun_out/${TAG}_bench.err; exit 1; }
cat gpurun_out/${TAG}_bench.json
timeout 300 python tools/profile_step.py 512 > gpurun_out/${TAG}_breakdown_b512.txt 2>&1
CMD="python bench.py --steps 1 --warmup 3 --no-extras --no-graph"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:${KRE} -s ${SKIP:-100} -c 3 \
    -o gpurun_out/${TAG}_${KRE}_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo done
