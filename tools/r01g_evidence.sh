#!/bin/bash
# Round-1 final evidence (run under gpurun from the repo root): bash tools/r01g_evidence.sh
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py 2>$o/r01g_bench.err > $o/r01g_bench.json; python tools/bench_brief.py < $o/r01g_bench.json
python bench.py --impl reference --steps 2 --warmup 1 2>$o/r01g_ref.err > $o/r01g_bench_reference.json; cut -c1-200 $o/r01g_bench_reference.json
python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 20 --warmup 3 --no-warp 2>$o/c3.err > $o/r01g_bench_config3.json; python tools/bench_brief.py < $o/r01g_bench_config3.json
CILQR_GRAPH=0 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $o/r01g_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $o/ncu_g.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $o/r01g_launches_config3.csv \
    python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 1 --warmup 1 --no-cpu --no-e2e --no-warp > $o/ncu_g3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ls_candidates -s 24 -c 1 -o $o/r01g_ls_candidates -f \
    python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 1 --warmup 0 --no-cpu --no-e2e --no-warp > $o/ncu_g4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ls_merit -s 25 -c 1 -o $o/r01g_ls_merit -f \
    python bench.py --batch 4096 --n-obs 10 --n-alphas 8 --steps 1 --warmup 0 --no-cpu --no-e2e --no-warp > $o/ncu_g5.log 2>&1
ls -la $o/*.ncu-rep | tail -3
