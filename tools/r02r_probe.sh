#!/bin/bash
# sweep kernel: resident warps per SM against L2 residency of a warp's records between its backward and forward pass
set -u
mkdir -p gpurun_out
B="--steps 20 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e"
for occ in 11 8 6 5 4 3; do
  echo "sweep occ $occ: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_SWEEP_OCC=$occ timeout 300 python bench.py $B 2>gpurun_out/r02r.err | python tools/bench_brief.py)"
done
for occ in 1; do
  echo "deriv occ $occ: $(CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_DERIV_OCC=$occ timeout 300 python bench.py $B 2>gpurun_out/r02r.err | python tools/bench_brief.py)"
done
for occ in 6 4; do
  CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_SWEEP_OCC=$occ PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02r_timeline_occ$occ.txt 2>&1
  grep sweep gpurun_out/r02r_timeline_occ$occ.txt | head -4
done
echo "--- e2e over 48 steps (product library)"
timeout 600 python bench.py --steps 50 --warmup 5 --no-warp --no-extra --no-cpu > gpurun_out/r02r_e2e.json 2> gpurun_out/r02r_e2e.err
python tools/bench_brief.py < gpurun_out/r02r_e2e.json
