"""Instruction mix and stall samples by opcode from an `ncu --set full --import-source on` report.

    ncu -i REPORT.ncu-rep --page source --print-source cuda,sass --csv > src.csv ; python tools/ncu_source_mix.py src.csv
"""
import collections
import csv
import sys


def num(x):
    try:
        return int(float(x.replace(",", "")))
    except Exception:
        return 0


rows = list(csv.reader(open(sys.argv[1])))
hdr, out = None, []
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        out.append(dict(zip(hdr, r)))
keys = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
# two views in one file: CUDA-C lines first, then SASS (rows whose second `Source` column holds an opcode and have an Address)
sass = [d for d in out if d.get("Address") not in (None, "", "-")]
tot = sum(num(d["# Samples"]) for d in sass)
by, instr, stall = collections.Counter(), collections.Counter(), collections.Counter()
for d in sass:
    src = d["Source"].split()
    if not src:
        continue
    op = (src[1] if src[0].startswith("@") and len(src) > 1 else src[0]).split(".")[0]
    by[op] += num(d["# Samples"])
    instr[op] += num(d["Instructions Executed"])
    for k in keys:
        stall[k] += num(d[k])
ti = sum(instr.values())
print("samples %d, warp instructions %d" % (tot, ti))
print("%-10s %9s %7s %12s %7s" % ("opcode", "samples", "share", "warp instr", "share"))
for op, n in by.most_common(16):
    print("%-10s %9d %6.1f%% %12d %6.1f%%" % (op, n, 100.0 * n / max(tot, 1), instr[op], 100.0 * instr[op] / max(ti, 1)))
print("stalls:", ", ".join("%s %d" % (k[6:], v) for k, v in stall.most_common(8)))
