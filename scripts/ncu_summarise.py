#!/usr/bin/env python
"""Summarise ncu raw-page CSVs (ncu -i X.ncu-rep --page raw --csv) into profiles/<tag>_ncu_summary.json and
profiles/ncu_summary.json (what bench.py's roofline.traffic reads).

    python scripts/ncu_summarise.py r1o gpurun_out/r1o_dp5_apply_raw.csv gpurun_out/r1o_error_sums_raw.csv

Round 2 captures hold ALL launches of one DP5 step attempt (6 x [pass_first<comb>, pass_mid, pass_centre, pass_mid,
pass_last], four of the last/first pairs fused into pass_last_first; the embedded error estimate rides on the last stage): dram_bytes_per_step is their sum.  Round-1 style
captures (one H-apply + the error sweep) are scaled: 6 * (mean first_comb + 2 mid + centre + last) + error_sums.
"""
import csv
import json
import os
import re
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "gpc__cycles_elapsed.avg.per_second",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TO_MS = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def short(name):
    m = re.search(r"(pass_\w+|error_sums_kernel)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in data:
        d = {"kernel": short(r[idx["Kernel Name"]])}
        for k in KEEP:
            if k in idx:
                try:
                    v = float(r[idx[k]].replace(",", ""))
                except ValueError:
                    continue
                u = units[idx[k]]
                if u in TO_BYTES:
                    v *= TO_BYTES[u]
                    u = "byte"
                elif k == "gpu__time_duration.sum":
                    v *= TO_MS.get(u, 1.0)
                    u = "ms"
                d[k] = v
                d.setdefault("_units", {})[k] = u
        out.append(d)
    return out


def main():
    tag, paths = sys.argv[1], sys.argv[2:]
    launches = [d for p in paths for d in load(p)]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def dram(d):
        return d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"]

    by = {}
    for d in launches:
        key = re.sub(r"<.*", "", d["kernel"])
        by.setdefault(key, []).append(d)
    mean = lambda xs: sum(xs) / len(xs)
    per_apply = (mean([dram(d) for d in by["pass_first"]]) + 2 * mean([dram(d) for d in by["pass_mid"]])
                 + mean([dram(d) for d in by["pass_centre"]]) + mean([dram(d) for d in by["pass_last"]]))
    err = mean([dram(d) for d in by.get("error_sums_kernel", [])]) if by.get("error_sums_kernel") else 0.0
    ms_apply = (mean([d["gpu__time_duration.sum"] for d in by["pass_first"]]) + 2 * mean([d["gpu__time_duration.sum"] for d in by["pass_mid"]])
                + mean([d["gpu__time_duration.sum"] for d in by["pass_centre"]]) + mean([d["gpu__time_duration.sum"] for d in by["pass_last"]]))
    n_fused = len(by.get("pass_last_first", []))   # a stage's pass_last fused with the next stage's pass_first
    whole_step = len(by["pass_last"]) + n_fused == 6 and len(by["pass_first"]) + n_fused == 6
    per_step = sum(dram(d) for d in launches) if whole_step else 6 * per_apply + err
    ms_step = sum(d["gpu__time_duration.sum"] for d in launches) if whole_step else None
    shares = {}
    if whole_step:
        for key, ds in by.items():
            shares[key] = {"launches": len(ds), "ncu_ms": sum(d["gpu__time_duration.sum"] for d in ds),
                           "share_of_step": sum(d["gpu__time_duration.sum"] for d in ds) / ms_step,
                           "dram_bytes": sum(dram(d) for d in ds)}
    summary = {
        "tag": tag,
        "source": "ncu --set full --clock-control none, python bench.py --steps 1 --warmup 3 --no-cpu --no-sharded (N=28, DP5); "
                  "round 2: all launches of the timed step attempt",
        "dram_bytes_per_step": per_step,
        "whole_step_capture": whole_step, "ncu_ms_per_step": ms_step, "kernel_shares": shares,
        "dram_bytes_per_happly": per_apply,
        "dram_bytes_error_sweep": err,
        "ncu_ms_per_happly": ms_apply,
        "note": "per-launch ncu times are cold-cache and serialised; bench.py's event timings are the rates",
        "launches": [{k: v for k, v in d.items() if k != "_units"} for d in launches],
        "units": launches[0].get("_units", {}) if launches else {},
    }
    for name in (f"{tag}_ncu_summary.json", "ncu_summary.json"):
        with open(os.path.join(root, "profiles", name), "w") as f:
            json.dump(summary, f, indent=1)
    S = 16.0 * (1 << 28)
    print(f"per H-apply: {per_apply / S:.2f} S DRAM, {ms_apply:.2f} ms under ncu; error sweep {err / S:.2f} S; per step {summary['dram_bytes_per_step'] / 1e9:.1f} GB")
    for d in launches:
        print(f"  {d['kernel'][:34]:34s} {d['gpu__time_duration.sum']:.3f} ms  R {d['dram__bytes_read.sum'] / S:.2f} S  W {d['dram__bytes_write.sum'] / S:.2f} S  "
              f"dram {d.get('dram__throughput.avg.pct_of_peak_sustained_elapsed', 0):.0f}%  fp64 {d.get('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 0):.0f}%  "
              f"regs {d.get('launch__registers_per_thread', 0):.0f}  long_sb {d.get('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 0):.1f}")


if __name__ == "__main__":
    main()
