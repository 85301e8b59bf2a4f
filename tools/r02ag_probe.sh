#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_solver.py tests/test_dropin.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "65536: $(timeout 300 python bench.py $B 2>gpurun_out/r02ag.err | python tools/bench_brief.py)"
echo "8192 fused: $(timeout 300 python bench.py $B --batch 8192 2>gpurun_out/r02ag.err | python tools/bench_brief.py)"
echo "config3: $(timeout 300 python bench.py --steps 20 --warmup 3 --no-warp --no-extra --no-cpu --no-e2e --batch 4096 --n-obs 10 --n-alphas 8 2>gpurun_out/r02ag.err | python tools/bench_brief.py)"
