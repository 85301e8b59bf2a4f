#!/bin/bash
# NOTE: probe of an experiment that was reverted (merged cost + derivative kernel, bench flag --split-eval): kept as the record of
# how profiles/r02p_* were taken; it does not run against the current tree.
# merged cost+derivative kernel (batch_eval_kernel): parity first, then A/B against the two-kernel rounds
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_solver.py -m gpu -x -q -k "fused_kernel_gives or solve_trace or batched_solve or golden or chunked or full_size or other_horizons" 2>&1 | tail -5
for mode in merged split; do
  flag=""; [ $mode = split ] && flag="--split-eval"
  timeout 300 python bench.py --steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e $flag > gpurun_out/r02p_$mode.json 2> gpurun_out/r02p_$mode.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r02p_$mode.json").read().strip().splitlines()[-1])
print("$mode", "value %.3f M/s  ms/step %.3f  solver_ms %.3f launches %s" % (d["value"]/1e6, d["ms_per_step"], d.get("roofline", {}).get("kernel_ms", float("nan")), d.get("gpu_launches")))
PY
done
for mode in merged split; do
  flag=""; [ $mode = split ] && flag="--split-eval"
  timeout 300 python bench.py --batch 8192 --fused 0 --steps 30 --warmup 5 --no-warp --no-extra --no-cpu --no-e2e $flag > gpurun_out/r02p_8192_$mode.json 2> gpurun_out/r02p_8192_$mode.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r02p_8192_$mode.json").read().strip().splitlines()[-1])
print("8192 rounds $mode", "value %.3f M/s  ms/step %.3f  solver_ms %.3f" % (d["value"]/1e6, d["ms_per_step"], d.get("roofline", {}).get("kernel_ms", float("nan"))))
PY
done
# timeline of one traced solve (merged)
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev 1 > gpurun_out/r02p_timeline_merged.txt 2>&1 || true
PYTHONPATH=. timeout 300 python tools/solve_timeline.py 65536 20 dev 0 > gpurun_out/r02p_timeline_split.txt 2>&1 || true
head -3 gpurun_out/r02p_timeline_merged.txt gpurun_out/r02p_timeline_split.txt
