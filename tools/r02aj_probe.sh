#!/bin/bash
# timing probe: derivative kernel with every scenario reading the same obstacle constants (L1 hits) — how much of round 0 is their latency?
set -u
mkdir -p gpurun_out
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02aj_a.txt 2>&1; grep "round  0" gpurun_out/r02aj_a.txt
CILQR_B200_LIB=build/libcilqr_sameobs.so PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev > gpurun_out/r02aj_b.txt 2>&1; grep "round  0" gpurun_out/r02aj_b.txt
