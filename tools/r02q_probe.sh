#!/bin/bash
# compute token chained on the device (no host wait between two handles' solves): parity of the threaded paths, then e2e
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_solver.py tests/test_dropin.py -m gpu -x -q -k "two_host_threads or device_resident_and_host or chunked or gather" 2>&1 | tail -4
for thr in 4 3 6; do
  timeout 400 python bench.py --steps 20 --warmup 5 --no-warp --no-extra --no-cpu --e2e-threads $thr > gpurun_out/r02q_e2e_$thr.json 2> gpurun_out/r02q_e2e_$thr.err
  echo "threads $thr: $(python tools/bench_brief.py < gpurun_out/r02q_e2e_$thr.json)"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r02q_e2e_$thr.json").read().strip().splitlines()[-1])
e=d["e2e"]; print("   e2e %.3f  u_only %.3f  one_thread %.3f  ceiling %.3f" % (e["value"]/1e6, e.get("u_only_value",0)/1e6, e.get("one_thread_value",0)/1e6, e["host_ceiling"]["value"]/1e6))
PY
done
