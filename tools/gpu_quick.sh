#!/bin/bash
# bench line (brief) + ncu launch list of one step; run under gpurun from the repo root:  bash tools/gpu_quick.sh <tag>
tag=${1:-x}
python bench.py --steps 10 --warmup 3 2>/dev/null | tee gpurun_out/bench_$tag.json | python tools/bench_brief.py
CILQR_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-warp > gpurun_out/ncu_$tag.log 2>&1
python tools/step_launches.py gpurun_out/launches_$tag.csv > gpurun_out/step_$tag.txt
grep sweep gpurun_out/step_$tag.txt | head -21 | awk '{printf "%s ", $2} END {print ""}'
tail -8 gpurun_out/step_$tag.txt
