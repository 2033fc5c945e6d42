"""Turn `ncu --set full` captures (gpurun_out/*.ncu-rep, read here with `ncu -i ... --page raw --csv`) into the
evidence files under profiles/:

    python tools/ncu_traffic.py <rep> <kernel label> <n> <gpus> <summary.txt> [<rep> ...]

For every capture: a text summary of the counters the roofline discussion uses (duration, FP64-pipe activity,
issue statistics, stall reasons per issue, occupancy, DRAM bytes) and one entry of profiles/r2_traffic.json
(DRAM bytes per launch, the mean over the captured launches), which bench.py reads for `roofline.traffic`.
Needs only the ncu command-line tool, no GPU."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fp64.sum",
    "sm__cycles_elapsed.max", "sm__cycles_active.min", "sm__cycles_active.avg", "sm__cycles_active.max", "smsp__cycles_active.avg",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def read_rep(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    return hdr, units, rows[2:]


def main():
    args = sys.argv[1:]
    if len(args) % 5:
        raise SystemExit(__doc__)
    table_path = os.path.join(ROOT, "profiles", "r2_traffic.json")
    table = {"what": "DRAM bytes per launch of the force kernel, mean over the launches of one `ncu --set full "
                     "--clock-control none` capture (dram__bytes_read.sum, dram__bytes_write.sum); written by tools/ncu_traffic.py",
             "captures": []}
    if os.path.exists(table_path):
        table = json.load(open(table_path))
    for k in range(0, len(args), 5):
        rep, label, n, gpus, summary = args[k], args[k + 1], int(args[k + 2]), int(args[k + 3]), args[k + 4]
        hdr, units, launches = read_rep(rep)
        col = {name: hdr.index(name) for name in WANT if name in hdr}
        name_col = hdr.index("Kernel Name")
        lines = ["# %s: ncu --set full --clock-control none --import-source on (%d launch(es) captured)" % (os.path.basename(rep), len(launches)),
                 "# kernel label: %s, N = %d, GPUs = %d" % (label, n, gpus)]
        rd, wr = [], []
        for li, row in enumerate(launches):
            lines.append("launch %d: %s" % (li, row[name_col]))
            for name in WANT:
                if name in col:
                    lines.append("  %-86s %16s %s" % (name, row[col[name]], units[col[name]]))
            for key, acc in (("dram__bytes_read.sum", rd), ("dram__bytes_write.sum", wr)):
                if key in col:
                    acc.append(float(row[col[key]].replace(",", "")) * SCALE.get(units[col[key]], 1.0))
        with open(os.path.join(ROOT, summary), "w") as f:
            f.write("\n".join(lines) + "\n")
        entry = {"kernel": label, "n": n, "gpus": gpus, "launches": len(launches),
                 "dram_bytes_read": sum(rd) / max(len(rd), 1), "dram_bytes_write": sum(wr) / max(len(wr), 1),
                 "source": summary}
        table["captures"] = [e for e in table["captures"] if not (e["kernel"] == label and e["n"] == n and e["gpus"] == gpus)] + [entry]
        print(entry)
    with open(table_path, "w") as f:
        json.dump(table, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
