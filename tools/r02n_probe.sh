#!/bin/bash
o=gpurun_out
python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "line_search or alpha or config3" -s 2>&1 | grep -v "^$" | tail -12
echo "--- config 3: fused / rounds"
for f in 1 0; do python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 20 --warmup 3 --no-warp --no-cpu --no-e2e --no-extra --fused $f 2>$o/n.err | python tools/bench_brief.py; tail -1 $o/n.err | cut -c1-200; done
echo "--- e2e with the token's kernels as a recorded graph"
python bench.py --steps 20 --warmup 3 --no-warp --no-cpu --no-extra 2>$o/n.err | python tools/bench_brief.py
