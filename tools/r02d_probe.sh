#!/bin/bash
# Round-2 second checkpoint: whole GPU suite, smoke, the new bench line.   bash tools/r02d_probe.sh
o=gpurun_out
python -m pytest tests -m gpu -x -q -s 2>&1 | grep -v "^$" | tail -25
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py 2>$o/r02d_bench.err > $o/r02d_bench.json; tail -3 $o/r02d_bench.err; python tools/bench_brief.py < $o/r02d_bench.json
python bench.py --batch 8192 --steps 20 --no-cpu --no-warp --no-e2e --no-extra 2>/dev/null | python tools/bench_brief.py
