#!/bin/bash
# NOTE: probe of an experiment that was reverted (tail split, bench flag --no-tail-split, option CILQR_OPT_TAIL_SPLIT): kept as the
# record of how profiles/r02ae_* were taken; it does not run against the current tree.
# tail split: parity, timing, timeline
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "tail_split or fused_kernel_gives or golden or full_size or chunked or trace or rank_shards" 2>&1 | tail -4
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "split:    $(timeout 300 python bench.py $B 2>gpurun_out/r02ae.err | python tools/bench_brief.py)"
echo "no split: $(timeout 300 python bench.py $B --no-tail-split 2>gpurun_out/r02ae.err | python tools/bench_brief.py)"
echo "split, no graph: $(timeout 300 python bench.py $B --no-graph 2>gpurun_out/r02ae.err | python tools/bench_brief.py)"
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02ae_timeline.txt 2>&1
head -2 gpurun_out/r02ae_timeline.txt; grep -E "round +(5|6|7|8|9) " gpurun_out/r02ae_timeline.txt | head -40; tail -5 gpurun_out/r02ae_timeline.txt
echo "e2e: $(timeout 600 python bench.py --steps 50 --warmup 5 --no-warp --no-extra --no-cpu 2>gpurun_out/r02ae.err | python tools/bench_brief.py)"
