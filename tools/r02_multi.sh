#!/bin/bash
# Driver-style multi-GPU bench line at N ranks (weak scaling headline + strong-scaling block + e2e), and the C++ gather demo.
N=${1:-2}
o=gpurun_out
./build/gather_demo $N 2>&1 | tail -1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps 20 --warmup 3 2>$o/r02_${N}gpu.err > $o/r02_bench_${N}gpu.json
tail -2 $o/r02_${N}gpu.err | cut -c1-300
python - <<PY
import json
d=json.load(open("$o/r02_bench_${N}gpu.json"))
e=d["e2e"]; s=d["strong_scaling"]
print("N=%d weak %.2f M/s ms/step %.3f | e2e %.2f M/s u_only %.2f ceiling %.2f | strong %.2f M/s ms/step %.3f solver_ms %.3f per gpu %d" % (d["n_gpus"], d["value"]/1e6, d["ms_per_step"], e["value"]/1e6, e["u_only_value"]/1e6, e["host_ceiling"]["value"]/1e6, s["value"]/1e6, s["ms_per_step"], s["solver_ms"], s["scenarios_per_gpu"]))
PY
