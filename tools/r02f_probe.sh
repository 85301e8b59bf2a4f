#!/bin/bash
# cost fused into init / sweep (round kernels), interleaved propagate.   bash tools/r02f_probe.sh
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
B="--steps 20 --warmup 3 --no-cpu --no-e2e --no-extra"
python bench.py $B 2>$o/f.err > $o/r02f_bench.json; tail -2 $o/f.err; python tools/bench_brief.py < $o/r02f_bench.json
python -c "
import json
d=json.load(open('gpurun_out/r02f_bench.json')); p=d['costmap_propagate']; print('propagate ms_per_map %.4f  fp64 frac %.4f' % (p['ms_per_map'], p['roofline']['frac']))"
for bb in 32768 16384 8192; do for f in 0 1; do echo "B=$bb fused=$f"; python bench.py $B --no-warp --batch $bb --fused $f 2>/dev/null | python tools/bench_brief.py; done; done
