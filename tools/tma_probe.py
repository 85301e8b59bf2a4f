"""Bring-up probe for the TMA-staged warp kernel: each case runs in its own process (a faulting kernel
poisons the CUDA context) and compares variant 2 (TMA) with variant 1 (direct).  Not a test, not the bench."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = {
    "A_fit128": (128, 128, 64, 64),
    "B_small80x60": (80, 60, 30, 22),
    "C_1024": (1024, 1024, 512, 512),
    "D_box_gt_rows": (64, 256, 64, 64),
}


def child(name):
    sys.path.insert(0, ROOT)
    import numpy as np
    import cilqr_b200
    from cilqr_b200 import scenarios
    rs, cs, rt, ct = CASES[name]
    G = cilqr_b200.CostmapGeom
    sg, tg = G(rs * 0.2, cs * 0.2, 0.2, 1.0, -2.0), G(rt * 0.2, ct * 0.2, 0.2, 0.0, 0.0)
    src = scenarios.make_source_map(rs, cs)
    poses = scenarios.make_poses(5, (1.0, -2.0), 1.0)
    v1, i1 = cilqr_b200.warp_batch(src, sg, tg, poses, variant=1)
    v2, i2 = cilqr_b200.warp_batch(src, sg, tg, poses, variant=2)
    print("%s: idx mismatches %d, value mismatches %d, oob %d" % (
        name, int(np.sum(i1 != i2)), int(np.sum(v1.view(np.uint32) != v2.view(np.uint32))), int(np.sum(i1 < 0))), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        child(sys.argv[1])
    else:
        for dbg in ("0", "1", "2"):
            for name in CASES:
                env = dict(os.environ, CILQR_TMA_DEBUG=dbg)
                r = subprocess.run([sys.executable, os.path.abspath(__file__), name], env=env, capture_output=True, text=True, timeout=300)
                tail = (r.stdout.strip().splitlines() or [""])[-1]
                err = [ln for ln in r.stderr.splitlines() if "Error" in ln or "error" in ln][-1:]
                print("dbg=%s rc=%d %s %s" % (dbg, r.returncode, tail, err), flush=True)
