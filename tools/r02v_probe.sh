#!/bin/bash
set -u
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
for c in 0 8 4 2 1; do
  echo "ib2 fork ctas $c: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_IB2_CTAS=$c timeout 300 python bench.py $B 2>gpurun_out/r02v.err | python tools/bench_brief.py)"
done
echo "e2e no fork: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_IB2_CTAS=0 timeout 600 python bench.py --steps 50 --warmup 5 --no-warp --no-extra --no-cpu 2>gpurun_out/r02v.err | python tools/bench_brief.py)"
