#!/bin/bash
# Round-2 closing evidence on the final code (run under gpurun from the repo root): bash tools/r02_final.sh
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py 2>$o/r02_bench.err > $o/r02_bench.json; python tools/bench_brief.py < $o/r02_bench.json
python bench.py --impl reference --steps 2 --warmup 1 2>$o/r02_ref.err > $o/r02_bench_reference.json; cut -c1-160 $o/r02_bench_reference.json
python bench.py --batch 8192 --steps 50 --no-warp --no-extra --no-cpu 2>/dev/null > $o/r02_bench_8192.json; python tools/bench_brief.py < $o/r02_bench_8192.json
PYTHONPATH=. python tools/solve_timeline.py 65536 20 dev > $o/r02_timeline.txt 2>&1; head -2 $o/r02_timeline.txt
# launch list of the default bench (direct launches so that ncu sees the round kernels one by one), time + DRAM bytes
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file $o/r02_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph > $o/ncu_r02_l.log 2>&1
# full sections: the three round kernels in round 0, the fused on-chip solve (8 192 scenarios), the warp at T = 4096
ncu --set full --clock-control none --import-source on -k regex:batch_derivs -s 20 -c 1 -o $o/r02_derivs -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:batch_sweep -s 20 -c 1 -o $o/r02_sweep -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:batch_cost -s 21 -c 1 -o $o/r02_cost -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02_c.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fused_solve -s 2 -c 1 -o $o/r02_fused -f \
    python bench.py --batch 8192 --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-warp > $o/ncu_r02_d.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:warp_tma -s 140 -c 1 -o $o/r02_warp_4096 -f \
    python bench.py --batch 4096 --steps 1 --warmup 0 --no-cpu --no-e2e --no-extra --warp-T 4096 > $o/ncu_r02_e.log 2>&1
ls -la $o/r02_*.ncu-rep $o/r02_launches.csv
