#!/bin/bash
set -u
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
for k in 1 2 3 4; do
echo "chunks $k: $(timeout 300 python bench.py $B --chunks $k 2>gpurun_out/r02ab.err | python tools/bench_brief.py)"
done
