"""Summarise an ncu report (raw page) into the handful of numbers the roofline argument needs."""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum.per_second",
        "sm__cycles_active.avg", "sm__cycles_active.min", "sm__cycles_active.max", "sm__cycles_elapsed.max",
        "sm__warps_active.avg.per_cycle_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]


def rows_of(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    return rows[0], rows[1], rows[2:]


def main(path):
    hdr, units, rows = rows_of(path)
    for r in rows:
        print("## kernel:", r[hdr.index("Kernel Name")][:110])
        for k in KEYS:
            if k in hdr:
                print("%-90s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))


def to_bytes(value, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return float(value.replace(",", "")) * mult


def traffic(name, path, out="profiles/traffic.json"):
    """python tools/ncu_summary.py --traffic <workload name> <report.ncu-rep>: record dram__bytes_read.sum +
    dram__bytes_write.sum of the report's (first) kernel launch in profiles/traffic.json, which bench.py reads for
    roofline.traffic at run time."""
    import json
    import os

    hdr, units, rows = rows_of(path)
    r = rows[0]
    rd = to_bytes(r[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_read.sum")])
    wr = to_bytes(r[hdr.index("dram__bytes_write.sum")], units[hdr.index("dram__bytes_write.sum")])
    head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    data = json.load(open(out)) if os.path.exists(out) else {}
    data[name] = {"dram_bytes": int(rd + wr), "read": int(rd), "write": int(wr),
                  "kernel": r[hdr.index("Kernel Name")][:100],
                  "source": "ncu --set full --clock-control none, %s (recorded at commit %s)" % (os.path.basename(path), head)}
    json.dump(data, open(out, "w"), indent=1, sort_keys=True)
    print(name, data[name])


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":
        traffic(sys.argv[2], sys.argv[3])
    else:
        main(sys.argv[1])
