#!/bin/bash
# launch list (time + DRAM bytes) of one step with the round kernels, after the record split
o=gpurun_out
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $o/r02x_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02x_l.log 2>&1
ls -la $o/r02x_launches.csv
ncu --set full --clock-control none --import-source on -k regex:batch_derivs -s 20 -c 1 -o $o/r02x_derivs -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02x_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:batch_sweep -s 20 -c 1 -o $o/r02x_sweep -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02x_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:batch_cost -s 21 -c 1 -o $o/r02x_cost -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --no-graph --no-warp > $o/ncu_r02x_c.log 2>&1
ls -la $o/r02x_*.ncu-rep
