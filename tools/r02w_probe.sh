#!/bin/bash
# sweep with 32-bit piece offsets (product); derivative kernel with 1 / 3 / 4 obstacles side by side
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "fused_kernel_gives or golden or batched_solve or chunked or line_search" 2>&1 | tail -3
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "product: $(timeout 300 python bench.py $B 2>gpurun_out/r02w.err | python tools/bench_brief.py)"
for u in 1 3 4; do
  echo "unroll $u: $(CILQR_B200_LIB=build/libcilqr_u$u.so timeout 300 python bench.py $B 2>gpurun_out/r02w.err | python tools/bench_brief.py)"
done
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02w_timeline.txt 2>&1
awk '/round/ {k=$5; v=$NF; gsub(/[()]/,"",v); s[k]+=v; r=$4; if (r<=2) a[k]+=v; else if (r<=7) b[k]+=v; else c[k]+=v} END {for (k in s) printf "%-8s total %7.1f  rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", k, s[k], a[k], b[k], c[k]}' gpurun_out/r02w_timeline.txt
