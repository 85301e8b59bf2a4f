#!/bin/bash
# launch list of one config-3 step: bash tools/ls_probe.sh <tag>
tag=${1:-x}
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 1 --warmup 1 --no-cpu --no-e2e --no-warp > gpurun_out/ncu_$tag.log 2>&1
python tools/step_launches.py gpurun_out/launches_$tag.csv > gpurun_out/step_$tag.txt; sed -n 8,15p gpurun_out/step_$tag.txt; tail -11 gpurun_out/step_$tag.txt
