#!/bin/bash
# the U* exchange: NCCL all-gather kernel against peer-to-peer copies on the copy engines.   bash tools/r02l_multi.sh N
N=${1:-2}
o=gpurun_out
for g in nccl peer; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --gather $g 2>$o/l_$g.err > $o/r02l_${N}gpu_$g.json
tail -2 $o/l_$g.err | cut -c1-300
python - <<PY
import json
try:
    d=json.load(open("$o/r02l_${N}gpu_$g.json"))
    s=d["strong_scaling"]
    print("gather $g: N=%d weak %.2f M/s ms/step %.3f | strong %.2f M/s ms/step %.3f solver_ms %.3f" % (d["n_gpus"], d["value"]/1e6, d["ms_per_step"], s["value"]/1e6, s["ms_per_step"], s["solver_ms"]))
except Exception as e: print("no line:", e)
PY
done
