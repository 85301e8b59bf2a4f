#!/bin/bash
# warp_tma_kernel: L2 prefetch (TMA) of the source box of a tile that starts kWarpAhead CTAs later
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_costmap.py tests/test_gpu_costmap_io.py -m gpu -x -q 2>&1 | tail -3
for T in 4096 2048 256; do
for a in product 0 740 2960; do
  lib=""; [ $a != product ] && lib="build/libcilqr_wa$a.so"
  CILQR_B200_LIB=$lib timeout 300 python bench.py --batch 4096 --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --warp-T $T > gpurun_out/r02ac_${T}_$a.json 2>gpurun_out/r02ac.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r02ac_${T}_$a.json").read().strip().splitlines()[-1])
w=d["costmap_warp"]
for s in w["sweep"]:
    print("T=%d ahead=%s: %.0f GB/s frac %.3f (value only %.0f)" % (s["T"], "$a", s["gbs"], s["frac"], s["value_only_gbs"]))
PY
done
done
