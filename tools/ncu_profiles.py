"""Summarise `ncu --set full` captures (gpurun_out/*.ncu-rep) into profiles/: one CSV row per kernel with the
metrics DESIGN.md quotes, and profiles/traffic.json (DRAM bytes per launch of the dominant kernel, read by bench.py).

    python tools/ncu_profiles.py r1i            # prefix of the captures in gpurun_out/
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}
WORKLOAD = {"poisson": "poisson_4096", "flow": "optical_flow_2048", "sfs": "sfs_2048", "arap": "arap_1m", "vol256": "volumetric_256", "vol512": "volumetric_512"}


def main():
    prefix = sys.argv[1]
    out_rows, traffic = [], {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    for fn in sorted(os.listdir(os.path.join(ROOT, "gpurun_out"))):
        if not (fn.startswith(prefix + "_") and fn.endswith("_full.ncu-rep")):
            continue
        name = fn[len(prefix) + 1:-len("_full.ncu-rep")]
        raw = subprocess.run(["ncu", "-i", os.path.join(ROOT, "gpurun_out", fn), "--page", "raw", "--csv"],
                             capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, vals = rows[0], rows[1], rows[2]
        d, u = dict(zip(hdr, vals)), dict(zip(hdr, units))
        rec = {"capture": fn, "kernel": d["Kernel Name"].split("(")[0]}
        for k in KEYS:
            if k in d:
                rec[k] = d[k] + (" " + u[k] if u[k] else "")
        def num(k):
            return float(d[k]) * UNIT.get(u[k], 1.0)
        byt = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
        rec["dram_bytes_per_launch"] = byt
        rec["dram_GBps"] = byt / num("gpu__time_duration.sum") / 1e9
        out_rows.append(rec)
        if name in WORKLOAD:
            traffic[WORKLOAD[name]] = {"bytes_per_launch": byt, "capture": "profiles/%s_ncu_full_summary.csv (%s)" % (prefix, fn)}
    cols = ["capture", "kernel", "dram_bytes_per_launch", "dram_GBps"] + KEYS
    with open(os.path.join(ROOT, "profiles", prefix + "_ncu_full_summary.csv"), "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=cols)
        w.writeheader()
        for r in out_rows:
            w.writerow({c: r.get(c, "") for c in cols})
    json.dump(traffic, open(tpath, "w"), indent=1, sort_keys=True)
    for r in out_rows:
        print("%-28s %-40s %8.1f us  %7.1f MB  %6.0f GB/s  issue %s  regs %s" % (
            r["capture"], r["kernel"][-40:], 1e6 * float(r["dram_bytes_per_launch"]) / (r["dram_GBps"] * 1e9),
            r["dram_bytes_per_launch"] / 1e6, r["dram_GBps"], r.get(KEYS[4], "?"), r.get(KEYS[6], "?")))


if __name__ == "__main__":
    main()
