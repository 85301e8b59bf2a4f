"""Per-launch device times of ONE bench step from an ncu launch list (gpu__time_duration.sum csv).

    python tools/step_launches.py gpurun_out/launches.csv [first_kernel_prefix]

Prints the launches of the last complete step (from the last occurrence of the obstacle-prep kernel on) in order."""
import csv
import io
import sys


def main():
    path = sys.argv[1]
    first = sys.argv[2] if len(sys.argv) > 2 else "prep_obstacles_cv_kernel"
    rows = list(csv.reader(io.StringIO("".join(l for l in open(path) if l.startswith('"')))))
    h = rows[0]
    ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
    mi = h.index("Metric Name")
    seq = [(r[ki].split("(")[0].replace("void ", "").replace("cilqr::", ""), float(r[vi]) / 1e3, r[gi]) for r in rows[1:] if r[mi] == "gpu__time_duration.sum"]
    idx = [i for i, (k, _, _) in enumerate(seq) if k.startswith(first)]
    start = idx[-1] if idx else 0
    tot, agg = 0.0, {}
    for k, v, g in seq[start:]:
        if k.startswith("fp64_peak"):
            break
        tot += v
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
        print("%-34s %9.1f us  %s" % (k[:34], v, g))
    print("-- step total %.1f us" % tot)
    for k, (n, v) in agg.items():
        print("%-34s n=%3d %9.1f us %5.1f%%" % (k[:34], n, v, 100 * v / tot))


if __name__ == "__main__":
    main()
