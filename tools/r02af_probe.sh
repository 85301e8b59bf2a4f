#!/bin/bash
# fused on-chip kernel against the round kernels by batch size (threshold of the automatic choice)
set -u
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
for b in 8192 12288 16384 20480 24576 32768; do
  for f in 1 0; do
    echo "batch $b fused=$f: $(timeout 300 python bench.py $B --batch $b --fused $f 2>gpurun_out/r02af.err | python tools/bench_brief.py)"
  done
done
