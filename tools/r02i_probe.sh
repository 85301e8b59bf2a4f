#!/bin/bash
o=gpurun_out
python -m pytest tests/test_gpu_costmap_io.py tests/test_gpu_costmap.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --batch 4096 2>$o/i.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); w=d['costmap_warp']
for e in w['sweep']: print('T %4d  %.0f GB/s  frac %.3f  value-only %.0f' % (e['T'], e['gbs'], e['frac'], e['value_only_gbs']))
for e in w['occupancy_in_out']['sweep']: print('int8 T %4d  ms/pose %.5f  %.0f GB/s  %.2f Gcells/s' % (e['T'], e['ms_per_pose'], e['gbs'], e['cells_per_s']/1e9))
print(w['occupancy_in_out']['float_path_ms_per_pose'])"
tail -3 $o/i.err
