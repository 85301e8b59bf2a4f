#!/bin/bash
# Fused kernel v3: phase cycle counts (instrumented build), timings, one ncu capture.   bash tools/r02b_probe.sh
o=gpurun_out
python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "fused or golden or shard or horizons" 2>&1 | tail -12
B="--steps 10 --warmup 3 --no-cpu --no-warp --no-e2e"
echo "--- phase cycles, B=65536 / 8192 (instrumented build)"
CILQR_B200_LIB=build/libcilqr_prof.so python bench.py --steps 1 --warmup 0 --no-cpu --no-warp --no-e2e --fused-from 0 2>&1 | grep "fused cta" | head -4
CILQR_B200_LIB=build/libcilqr_prof.so python bench.py --steps 1 --warmup 0 --no-cpu --no-warp --no-e2e --fused-from 0 --batch 8192 2>&1 | grep "fused cta" | head -4
for ff in 0 2; do echo "B=65536 fused-from $ff"; python bench.py $B --fused-from $ff 2>$o/a.err | python tools/bench_brief.py; done
for g in 16 24 28; do echo "B=65536 fused-from 0 slots $g"; python bench.py $B --fused-from 0 --fused-slots $g 2>$o/a.err | python tools/bench_brief.py; done
echo "B=65536 threads 128 slots 13";  python bench.py $B --fused-from 0 --fused-slots 13 --fused-threads 128 2>$o/a.err | python tools/bench_brief.py
echo "B=65536 threads 192 slots 20";  python bench.py $B --fused-from 0 --fused-slots 20 --fused-threads 192 2>$o/a.err | python tools/bench_brief.py
for bb in 32768 16384 8192 4096; do echo "B=$bb fused-from 0 / 99"; python bench.py $B --batch $bb --fused-from 0 2>$o/a.err | python tools/bench_brief.py;  python bench.py $B --batch $bb --fused-from 99 2>$o/a.err | python tools/bench_brief.py; done
for g in 8 14; do echo "B=8192 slots $g";  python bench.py $B --batch 8192 --fused-from 0 --fused-slots $g 2>$o/a.err | python tools/bench_brief.py; done
ncu --set full --clock-control none --import-source on -k regex:fused_solve -c 1 -o $o/r02b_fused -f python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-warp --fused-from 0 > $o/ncu_r02b.log 2>&1
ls -la $o/r02b_fused.ncu-rep
