#!/bin/bash
o=gpurun_out
python -m pytest tests/test_propagate.py tests/test_dropin.py -m gpu -x -q -s 2>&1 | grep -v "^$" | tail -9
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:propagate -c 30 --csv --log-file $o/r02h_prop.csv python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-extra --batch 4096 --warp-T 256 > $o/ncu_r02h.log 2>&1
grep propagate $o/r02h_prop.csv | tail -4 | cut -d, -f5,13-
python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-extra --batch 4096 --warp-T 256 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); p=d['costmap_propagate']; print('propagate ms_per_map %.4f  fp64 frac %.4f' % (p['ms_per_map'], p['roofline']['frac']))"
