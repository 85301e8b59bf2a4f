#!/bin/bash
# costmap warp: parity tests + the config-5 sweep; run under gpurun: bash tools/warp_quick.sh
python -m pytest tests/test_gpu_costmap.py tests/test_gpu_costmap_io.py -m gpu -x -q 2>&1 | tail -2
for T in 256 512 1024 2048 4096; do
python bench.py --batch 4096 --steps 3 --warmup 3 --no-cpu --no-e2e --warp-T $T 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['costmap_warp']
print('%-8d %.1f GB/s   %.1f GB/s   %.3f   %s' % ($T, d['value'], d['value_only_gbs'], d['roofline']['frac'], d['config']['workload']))"
done
