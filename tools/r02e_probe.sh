#!/bin/bash
# propagate kernel check + timing, compute-sanitizer memcheck / racecheck over small-batch GPU tests.   bash tools/r02e_probe.sh
o=gpurun_out
python -m pytest tests/test_propagate.py tests/test_dropin.py -m gpu -x -q -s 2>&1 | grep -v "^$" | tail -12
python bench.py --steps 3 --warmup 1 --no-cpu --no-e2e --no-extra --batch 4096 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); p=d['costmap_propagate']; print('propagate ms_per_map %.4f  fp64 frac %.4f' % (p['ms_per_map'], p['roofline']['frac']))"
SEL='golden or (fused and 9-2) or (fused and 41-4) or line_search or horizons or edge_cases or two_live'
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --log-file $o/r02e_memcheck.log python -m pytest tests/test_gpu_solver.py tests/test_propagate.py tests/test_gpu_costmap.py -m gpu -q -x -k "$SEL or default_30x20 or ragged or config2" > $o/r02e_memcheck.out 2>&1; echo "memcheck rc=$?"; tail -2 $o/r02e_memcheck.out; tail -3 $o/r02e_memcheck.log
timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 9 --log-file $o/r02e_racecheck.log python -m pytest tests/test_gpu_solver.py tests/test_propagate.py -m gpu -q -x -k "n50_obs3 or (fused and 9-2) or line_search or default_30x20 or two_live" > $o/r02e_racecheck.out 2>&1; echo "racecheck rc=$?"; tail -2 $o/r02e_racecheck.out; tail -3 $o/r02e_racecheck.log
