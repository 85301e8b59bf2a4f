#!/bin/bash
# sweep ring depth (records in flight per lane)
set -u
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "ring 3 (product): $(timeout 300 python bench.py $B 2>gpurun_out/r02y.err | python tools/bench_brief.py)"
for r in 2 4 5; do
  echo "ring $r: $(CILQR_B200_LIB=build/libcilqr_ring$r.so timeout 300 python bench.py $B 2>gpurun_out/r02y.err | python tools/bench_brief.py)"
  CILQR_B200_LIB=build/libcilqr_ring$r.so PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02y_timeline_ring$r.txt 2>&1
  awk -v R=$r '/sweep/ {v=$NF; gsub(/[()]/,"",v); r=$4; if (r<=2) a+=v; else if (r<=7) b+=v; else c+=v} END {printf "   ring %d sweep rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", R, a, b, c}' gpurun_out/r02y_timeline_ring$r.txt
done
CILQR_B200_LIB=build/libcilqr_ring4.so timeout 600 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "fused_kernel_gives or golden or batched_solve" 2>&1 | tail -2
