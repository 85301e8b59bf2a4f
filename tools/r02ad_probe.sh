#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_costmap.py -m gpu -x -q 2>&1 | tail -2
run() { # T, label, env
  python - "$1" "$2" <<PY
import json,sys
d=json.loads(open("gpurun_out/r02ad.json").read().strip().splitlines()[-1])
for s in d["costmap_warp"]["sweep"]:
    print("T=%d %s: %.0f GB/s frac %.3f (value only %.0f)" % (s["T"], sys.argv[2], s["gbs"], s["frac"], s["value_only_gbs"]))
PY
}
for T in 4096 2048 256; do
  timeout 300 python bench.py --batch 4096 --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --warp-T $T > gpurun_out/r02ad.json 2>gpurun_out/r02ad.err; run $T product
done
for a in 0 148 296 444 592 740 888 1110; do
  CILQR_B200_LIB=build/libcilqr_probe.so CILQR_PROBE_WARP_AHEAD=$a timeout 300 python bench.py --batch 4096 --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra --warp-T 4096 > gpurun_out/r02ad.json 2>gpurun_out/r02ad.err; run 4096 "ahead=$a"
done
