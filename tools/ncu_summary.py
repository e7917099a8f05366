"""Turn ncu output brought back in gpurun_out/ into the small text summaries kept under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches_r1.csv  > profiles/r1_launches.txt
  python tools/ncu_summary.py report   gpurun_out/prof_r1_trace.ncu-rep > profiles/r1_k_trace.txt

`launches` reads the CSV of `ncu --metrics gpu__time_duration.sum --clock-control none --csv`
and prints per-kernel launch counts, total / mean time and the share of the captured step.
`report` reads a `--set full` .ncu-rep (through `ncu -i ... --page raw --csv`) and prints the
metrics DESIGN.md / bench.py quote.
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed.sum",
    "l1tex__t_sector_hit_rate.pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_lg_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio",
    "smsp__average_warp_latency_issue_stalled_wait.ratio",
    "smsp__average_warp_latency_issue_stalled_not_selected.ratio",
    "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_branch_resolving.ratio",
    "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
    "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum",
]


def short(name):
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"gdpt::", "", name)
    m = re.match(r"([A-Za-z0-9_]+(?:<[^>(]*>)?)", name)
    return m.group(1) if m else name[:60]


def launches(path):
    rows = [r for r in csv.reader(open(path, newline="")) if r]
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[hi]
    kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        t = float(r[mv].replace(",", ""))
        unit = r[mu]
        t_us = t * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        a = agg.setdefault(short(r[kn]), [0, 0.0])
        a[0] += 1
        a[1] += t_us
    total = sum(a[1] for a in agg.values())
    print("# source: %s (ncu --metrics gpu__time_duration.sum --clock-control none; serialised, cold-cache launches)" % path)
    print("%-44s %8s %12s %12s %7s" % ("kernel", "launches", "total_us", "mean_us", "share"))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-44s %8d %12.1f %12.1f %6.1f%%" % (k, n, t, t / n, 100 * t / total))
    print("%-44s %8d %12.1f" % ("total", sum(a[0] for a in agg.values()), total))


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h, units = rows[hi], rows[hi + 1]
    print("# source: %s (ncu --set full --clock-control none --import-source on)" % path)
    for r in rows[hi + 2:]:
        print("kernel: %s" % r[h.index("Kernel Name")])
        for m in METRICS:
            if m in h:
                i = h.index(m)
                print("  %-72s %s %s" % (m, r[i], units[i]))
        if "dram__bytes_read.sum" in h:
            print()


def as_json(path):
    """The handful of per-launch numbers bench.py quotes next to its live measurements."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h, units, r = rows[hi], rows[hi + 1], rows[hi + 2]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
    def val(m):
        i = h.index(m)
        return float(r[i].replace(",", "")) * scale.get(units[i], 1)
    d = {"source": path, "kernel": short(r[h.index("Kernel Name")]),
         "duration_s": val("gpu__time_duration.sum"),
         "dram_bytes": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
         "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
         "threads_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
         "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
         "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
         "registers": val("launch__registers_per_thread"), "grid": val("launch__grid_size")}
    print(json.dumps(d, indent=1))


if __name__ == "__main__":
    {"launches": launches, "report": report, "json": as_json}[sys.argv[1]](sys.argv[2])
