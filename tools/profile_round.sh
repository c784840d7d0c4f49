#!/bin/bash
# Refresh the profiler evidence under gpurun_out/ (copy what should be judged into profiles/).
# Usage (GPU box): bash tools/profile_round.sh <tag>
TAG=${1:-r01}
CMD="python bench.py --n 200000 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 200 $CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -s 6000 -c 4000 --csv \
    --log-file gpurun_out/${TAG}_launches_n200k.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:irls_step -s 400 -c 2 \
    -o gpurun_out/${TAG}_irls_step_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/ | tail -n 8
