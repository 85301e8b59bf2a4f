#!/bin/bash
# Round-2 first light of the fused kernel: bash tools/r02a_probe.sh   (under gpurun from the repo root)
o=gpurun_out
python -m pytest tests/test_gpu_solver.py -m gpu -x -q 2>&1 | tail -15
B="--steps 10 --warmup 3 --no-cpu --no-warp --no-e2e"
for ff in 99 0 2 3 4; do echo "B=65536 fused-from $ff"; python bench.py $B --fused-from $ff 2>$o/a.err | python tools/bench_brief.py; tail -2 $o/a.err | cut -c1-300; done
for g in 7 10 12; do echo "B=65536 fused-from 0 slots $g"; python bench.py $B --fused-from 0 --fused-slots $g 2>$o/a.err | python tools/bench_brief.py; done
for ff in 99 0; do echo "B=8192 fused-from $ff"; python bench.py $B --batch 8192 --fused-from $ff 2>$o/a.err | python tools/bench_brief.py; tail -2 $o/a.err | cut -c1-300; done
for g in 4 7 10; do echo "B=8192 fused-from 0 slots $g"; python bench.py $B --batch 8192 --fused-from 0 --fused-slots $g 2>$o/a.err | python tools/bench_brief.py; done
