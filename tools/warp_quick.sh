#!/bin/bash
# costmap warp: parity tests + the config-5 number at a few sizes; run under gpurun: bash tools/warp_quick.sh
python -m pytest tests/test_gpu_costmap.py tests/test_gpu_costmap_io.py -m gpu -x -q 2>&1 | tail -2
for T in 2048 1024 4096; do
python bench.py --batch 4096 --steps 3 --warmup 3 --no-cpu --no-e2e --warp-T $T 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['costmap_warp']
print('T', $T, 'value+index %.1f GB/s (frac %.3f)  value only %.1f GB/s' % (d['value'], d['roofline']['frac'], d['value_only_gbs']))"
done
