"""Print the few numbers of a bench.py JSON line that change from run to run (stdin -> stdout)."""
import json
import sys

for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    r = d.get("roofline", {})
    print(f"value {d['value'] / 1e6:.3f} M/s  ms/step {d['ms_per_step']:.3f}  solver_ms {r.get('kernel_ms', 0):.3f}  "
          f"e2e {((d.get('e2e') or {}).get('value') or 0) / 1e6:.3f} M/s (1 thread {((d.get('e2e') or {}).get('one_thread_value') or 0) / 1e6:.3f})  launches {d.get('gpu_launches')}  fp64 frac {(r.get('frac') or 0):.3f}  "
          f"clocks {d.get('clocks', {}).get('sm_mhz')}")
