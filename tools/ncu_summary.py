"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv          > profiles/rNN_launches.txt
    python tools/ncu_summary.py full gpurun_out/prof_solve.ncu-rep [...]  > profiles/rNN_solve_full.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__average_warp_latency_per_inst_issued.ratio",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_static",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_warps", "sm__maximum_warps_per_active_cycle_pct", "smsp__warps_eligible.avg.per_cycle_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def read_csv_lines(text):
    lines = [ln for ln in text.splitlines() if ln.startswith('"')]
    return list(csv.reader(io.StringIO("\n".join(lines))))


def launches(path):
    rows = read_csv_lines(open(path).read())
    h = rows[0]
    ki, vi, gi, bi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size"), h.index("Block Size")
    agg = collections.OrderedDict()
    mi = h.index("Metric Name")
    for r in rows[1:]:
        if r[mi] != "gpu__time_duration.sum":
            continue
        k = r[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0, r[gi], r[bi]])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    print("# per-kernel device time from `ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares)")
    print("%-44s %5s %12s %12s %7s  %s" % ("kernel", "n", "total_ms", "avg_ms", "share", "grid x block"))
    for k, a in agg.items():
        print("%-44s %5d %12.3f %12.3f %6.1f%%  %s x %s" % (k, a[0], a[1] / 1e6, a[1] / a[0] / 1e6, 100 * a[1] / tot, a[2], a[3]))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = read_csv_lines(out)
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        print("## %s   grid %s block %s" % (r[h.index("Kernel Name")][:100], r[h.index("Grid Size")], r[h.index("Block Size")]))
        for k in KEYS:
            if k in h:
                i = h.index(k)
                print("%-75s %18s %s" % (k, r[i], units[i]))
        stalls = []
        for i, n in enumerate(h):
            if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio") and "not_issued" not in n:
                try:
                    stalls.append((float(r[i].replace(",", "")), n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        print("stall reasons (warps per issue-active cycle): " + ", ".join("%s %.2f" % (n, v) for v, n in stalls[:8]))
        print()


def traffic(path):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per launch from a launch-list csv that carries those two
    metrics: the solver's kernels summed over the LAST complete bench step, the costmap warp kernel averaged per launch.
    Prints the JSON that bench.py reads from profiles/traffic.json."""
    import json
    rows = read_csv_lines(open(path).read())
    h = rows[0]
    ii, ki, mi, vi, ui = h.index("ID"), h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.OrderedDict()
    for r in rows[1:]:
        if not r[mi].startswith("dram__bytes"):
            continue
        k = per.setdefault(r[ii], [r[ki].split("(")[0].replace("void ", "").replace("cilqr::", ""), 0.0, r[h.index("Grid Size")]])
        k[1] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    seq = list(per.values())
    grids = [g for _, _, g in seq]
    seq = [(k, v) for k, v, _ in seq]
    starts = [i for i, (k, _) in enumerate(seq) if k.startswith("prep_obstacles_cv_kernel")]
    out = {}
    if starts:
        step = []
        for k, v in seq[starts[-1]:]:
            if not (k.startswith("prep_") or k.startswith("batch_")):
                break
            step.append((k, v))
        out["solver_dram_bytes_per_step"] = sum(v for _, v in step)
        out["solver_dram_bytes_by_kernel"] = {}
        for k, v in step:
            out["solver_dram_bytes_by_kernel"][k] = out["solver_dram_bytes_by_kernel"].get(k, 0.0) + v
    # costmap warp (float value + index map): per launch, by target edge T — the grid is ((T / 64)^2 tiles, poses)
    by_t = collections.OrderedDict()
    for (k, v), g in zip(seq, grids):
        if k.startswith("warp_tma_kernel") or k.startswith("warp_direct_kernel"):
            tiles = int(g.strip("()").split(",")[0])
            by_t.setdefault(64 * int(round(tiles ** 0.5)), []).append(v)
    for t, vs in by_t.items():
        vs = vs[len(vs) // 2:]  # the second half of a size's launches are the value + index passes (bench_warp's order)
        out["warp_dram_bytes_per_launch_T%d" % t] = sum(vs) / len(vs)
    fused = [v for k, v in seq if k.startswith("fused_solve_kernel")]
    if fused:
        out["fused_solve_dram_bytes_per_launch"] = sum(fused) / len(fused)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    mode = sys.argv[1]
    for p in sys.argv[2:]:
        {"launches": launches, "traffic": traffic}.get(mode, full)(p)
