#!/bin/bash
set -u
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "product: $(timeout 300 python bench.py $B 2>gpurun_out/r02aa.err | python tools/bench_brief.py)"
echo "L1 prefetch of obstacle constants: $(CILQR_B200_LIB=build/libcilqr_l1pf.so timeout 300 python bench.py $B 2>gpurun_out/r02aa.err | python tools/bench_brief.py)"
CILQR_B200_LIB=build/libcilqr_l1pf.so PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02aa_timeline.txt 2>&1
awk '/round/ {k=$5; v=$NF; gsub(/[()]/,"",v); s[k]+=v; r=$4; if (r<=2) a[k]+=v; else if (r<=7) b[k]+=v; else c[k]+=v} END {for (k in s) printf "%-8s total %7.1f  rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", k, s[k], a[k], b[k], c[k]}' gpurun_out/r02aa_timeline.txt
timeout 900 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "obstacle_set_is_complete" 2>&1 | tail -2
