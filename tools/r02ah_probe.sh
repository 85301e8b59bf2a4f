#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_solver.py tests/test_dropin.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "65536: $(timeout 300 python bench.py $B 2>gpurun_out/r02ah.err | python tools/bench_brief.py)"
echo "8192 fused: $(timeout 300 python bench.py $B --batch 8192 2>gpurun_out/r02ah.err | python tools/bench_brief.py)"
echo "config3: $(timeout 300 python bench.py --steps 20 --warmup 3 --no-warp --no-extra --no-cpu --no-e2e --batch 4096 --n-obs 10 --n-alphas 8 2>gpurun_out/r02ah.err | python tools/bench_brief.py)"
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02ah_timeline.txt 2>&1
awk '/round/ {k=$5; v=$NF; gsub(/[()]/,"",v); s[k]+=v; r=$4; if (r<=2) a[k]+=v; else if (r<=7) b[k]+=v; else c[k]+=v} END {for (k in s) printf "%-8s total %7.1f  rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", k, s[k], a[k], b[k], c[k]}' gpurun_out/r02ah_timeline.txt
