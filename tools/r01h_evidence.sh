#!/bin/bash
# Round-1 closing evidence (run under gpurun from the repo root): bash tools/r01h_evidence.sh
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py 2>$o/r01h_bench.err > $o/r01h_bench.json; python tools/bench_brief.py < $o/r01h_bench.json
python bench.py --impl reference --steps 2 --warmup 1 2>$o/r01h_ref.err > $o/r01h_bench_reference.json; cut -c1-160 $o/r01h_bench_reference.json
python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 20 --warmup 3 --no-warp 2>$o/c3.err > $o/r01h_bench_config3.json; python tools/bench_brief.py < $o/r01h_bench_config3.json
# launch list of the default bench (direct launches so that ncu sees the round kernels one by one)
CILQR_GRAPH=0 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $o/r01h_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $o/ncu_h_l.log 2>&1
# round 0 of the solve (65 536 scenarios): cost, derivative and sweep kernels, full sections
CILQR_GRAPH=0 ncu --set full --clock-control none --import-source on -k 'regex:batch_(cost|derivs|sweep)' -c 3 -o $o/r01h_round0 -f \
    python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-warp > $o/ncu_h_a.log 2>&1
# costmap warp (value + index map, 2048^2 from 4096^2, 64 poses) and the uncertainty-propagation kernel
ncu --set full --clock-control none --import-source on -k regex:warp_ -s 40 -c 1 -o $o/r01h_warp -f \
    python bench.py --batch 4096 --steps 1 --warmup 0 --no-cpu --no-e2e > $o/ncu_h_b.log 2>&1
ls -la $o/r01h_*.ncu-rep $o/r01h_launches.csv
