#!/bin/bash
# e2e knobs (host threads x chunks), call-by-call breakdown, ncu of the tail sweep; run under gpurun: bash tools/r01h_probe.sh
o=gpurun_out
PYTHONPATH=. python tools/e2e_breakdown.py 2>&1 | tail -12
for t in 2 3 4 6; do echo "threads $t"; python bench.py --e2e-threads $t --no-cpu --no-warp --steps 12 --warmup 3 2>/dev/null | python tools/bench_brief.py; done
for c in 2 8; do echo "chunks $c"; CILQR_CHUNKS=$c python bench.py --e2e-threads 3 --no-cpu --no-warp --steps 12 --warmup 3 2>/dev/null | python tools/bench_brief.py; done
CILQR_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:batch_sweep -s 12 -c 1 -o $o/r01h_sweep_tail -f \
    python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-warp > $o/ncu_h1.log 2>&1
ls -la $o/*.ncu-rep | tail -2
