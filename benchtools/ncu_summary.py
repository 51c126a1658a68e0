#!/usr/bin/env python
"""Reduce `ncu --page raw --csv` exports to the counters SURVEY.md 8(d) asks for.

    ncu -i prof.ncu-rep --page raw --csv > prof_raw.csv      (on the GPU box or here)
    python benchtools/ncu_summary.py label=prof_raw.csv ... --out profiles/r2_kernels_ncu.csv \
           --traffic profiles/traffic.json

One output row per profiled launch; `traffic.json` gets dram read + write bytes per launch and
workload label (what bench.py reports as roofline.traffic).  Tuning tool, not product code.
"""
from __future__ import annotations

import csv
import json
import sys

WANT = {
    "gpu__time_duration.sum": "time_us",
    "dram__bytes_read.sum": "dram_read_MB",
    "dram__bytes_write.sum": "dram_write_MB",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum": "ld_sectors",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum": "ld_requests",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "long_scoreboard_per_issue",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "sm__cycles_active.avg": "sm_cycles_active",
    "sm__cycles_elapsed.avg": "sm_cycles_elapsed",
}
SCALE = {"dram_read_MB": 1e-6, "dram_write_MB": 1e-6}


def _num(x):
    try:
        return float(str(x).replace(",", ""))
    except ValueError:
        return None


def _to_unit(value, unit, want):
    """ncu prints a unit per column; bring bytes to bytes and time to microseconds."""
    if value is None:
        return None
    u = (unit or "").strip().lower()
    if want in ("dram_read_MB", "dram_write_MB"):
        mult = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)
        return value * mult * 1e-6
    if want == "time_us":
        mult = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "second": 1e6}.get(u, 1e-3)
        return value * mult
    return value


def rows_of(path):
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = list(csv.reader(lines))
    if len(rd) < 3:
        return []
    head, units = rd[0], rd[1]
    out = []
    for r in rd[2:]:
        if len(r) != len(head):
            continue
        rec = {}
        for h, u, v in zip(head, units, r):
            if h == "Kernel Name":
                rec["kernel"] = v
            elif h in WANT:
                rec[WANT[h]] = _to_unit(_num(v), u, WANT[h])
        if "kernel" in rec:
            out.append(rec)
    return out


def main(argv):
    out_path = traffic_path = None
    items = []
    i = 0
    while i < len(argv):
        if argv[i] == "--out":
            out_path = argv[i + 1]
            i += 2
        elif argv[i] == "--traffic":
            traffic_path = argv[i + 1]
            i += 2
        else:
            label, path = argv[i].split("=", 1)
            items.append((label, path))
            i += 1
    cols = ["label", "kernel"] + list(dict.fromkeys(WANT.values())) + ["sectors_per_request", "dram_total_MB"]
    table, traffic = [], {}
    for label, path in items:
        per_label = 0.0
        for rec in rows_of(path):
            rec["label"] = label
            if rec.get("ld_sectors") and rec.get("ld_requests"):
                rec["sectors_per_request"] = rec["ld_sectors"] / rec["ld_requests"]
            rec["dram_total_MB"] = (rec.get("dram_read_MB") or 0.0) + (rec.get("dram_write_MB") or 0.0)
            per_label += rec["dram_total_MB"] * 1e6
            table.append(rec)
        if per_label:
            traffic[label] = per_label
    w = csv.writer(open(out_path, "w", newline="") if out_path else sys.stdout)
    w.writerow(cols)
    for rec in table:
        w.writerow([("%.4g" % rec[c]) if isinstance(rec.get(c), float) else rec.get(c, "") for c in cols])
    if traffic_path:
        with open(traffic_path, "w") as f:
            json.dump(traffic, f, indent=1)


if __name__ == "__main__":
    main(sys.argv[1:])
