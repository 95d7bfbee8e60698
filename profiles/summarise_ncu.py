#!/usr/bin/env python
"""Turn an `ncu --set full` report into the committed evidence:

    python profiles/summarise_ncu.py gpurun_out/r02c_full_n400.ncu-rep profiles/r02c_ncu_full_n400.csv

writes the per-launch summary CSV (DRAM bytes, DRAM throughput, duration, registers, occupancy, issue, stall) and
updates profiles/ncu_traffic.json: per bench timer (mules_iter, ...) the mean `dram__bytes_read.sum +
dram__bytes_write.sum` per launch, which bench.py reports as `roofline.traffic` beside the algorithmic bytes."""
import csv
import io
import json
import os
import subprocess
import sys

WANT = ["Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__time_duration.sum", "launch__block_size",
        "launch__grid_size", "launch__registers_per_thread", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__inst_executed.sum"]
TIMER_OF = [("grad_tensor", "grad_tensor"), ("div_tensor", "div_tensor"), ("mules_setup", "mules_setup"),
            ("mules_iter", "mules_iter"), ("mules_final", "mules_final"), ("mules_update", "mules_final"), ("limit_sum", "limit_sum"),
            ("flux_assembly", "flux_assembly"), ("pred_jacobi", "predictor"), ("alpha_transfer", "alpha_transfer")]


def to_bytes(v, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    return float(v) * scale


def main(rep, out_csv):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(WANT)
        w.writerow([units[idx[x]] for x in WANT])
        for r in data:
            w.writerow([r[idx[x]] for x in WANT])
    acc = {}
    for r in data:
        name = r[idx["Kernel Name"]]
        if "kv" not in name:           # volume kernels only (kv_*, kvp_*); the BND twins are O(surface)
            continue
        for key, timer in TIMER_OF:
            if key in name:
                b = (to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
                     + to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]]))
                acc.setdefault(timer, []).append(b)
                break
    here = os.path.dirname(os.path.abspath(__file__))
    out = {"source": os.path.basename(out_csv), "what": "mean dram__bytes_read.sum + dram__bytes_write.sum per launch, "
           "ncu --set full --clock-control none, bench.py configs[4] 400^3 (64 M cells)",
           "bytes_per_launch": {k: sum(v) / len(v) for k, v in acc.items()},
           "launches_captured": {k: len(v) for k, v in acc.items()}}
    with open(os.path.join(here, "ncu_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
