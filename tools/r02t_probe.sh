#!/bin/bash
# after the record split: cost kernel with next-group prefetch (product), sweep occupancy again, derivative kernel with 320 threads
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "fused_kernel_gives or golden or batched_solve or chunked" 2>&1 | tail -3
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
echo "product: $(timeout 300 python bench.py $B 2>gpurun_out/r02t.err | python tools/bench_brief.py)"
for occ in 8 6 4; do
  echo "sweep occ $occ: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_SWEEP_OCC=$occ timeout 300 python bench.py $B 2>gpurun_out/r02t.err | python tools/bench_brief.py)"
done
for occ in 2 1; do
  echo "cost occ $occ: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_COST_OCC=$occ timeout 300 python bench.py $B 2>gpurun_out/r02t.err | python tools/bench_brief.py)"
done
echo "derivs 320 threads: $(CILQR_B200_LIB=build/libcilqr_d320.so timeout 300 python bench.py $B 2>gpurun_out/r02t.err | python tools/bench_brief.py)"
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02t_timeline.txt 2>&1
awk '/round/ {k=$5; v=$NF; gsub(/[()]/,"",v); s[k]+=v; r=$4; if (r<=2) a[k]+=v; else if (r<=7) b[k]+=v; else c[k]+=v} END {for (k in s) printf "%-8s total %7.1f  rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", k, s[k], a[k], b[k], c[k]}' gpurun_out/r02t_timeline.txt
CILQR_B200_LIB=build/libcilqr_d320.so PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02t_timeline_d320.txt 2>&1
awk '/round/ {k=$5; v=$NF; gsub(/[()]/,"",v); s[k]+=v; r=$4; if (r<=2) a[k]+=v; else if (r<=7) b[k]+=v; else c[k]+=v} END {for (k in s) printf "d320 %-8s total %7.1f  rounds0-2 %7.1f  rounds3-7 %7.1f  rounds8+ %7.1f\n", k, s[k], a[k], b[k], c[k]}' gpurun_out/r02t_timeline_d320.txt
