#!/usr/bin/env python3
"""Turn ncu CSV exports into the summaries kept under profiles/.

  python tools/ncu_summary.py launches <launches.csv> <out.md> "<command>"
      launches.csv = `ncu --metrics gpu__time_duration.sum --clock-control none --csv ...` log
  python tools/ncu_summary.py full <raw.csv> <out.md> <traffic.json> <rows> "<command>"
      raw.csv = `ncu -i <rep> --page raw --csv` of an `ncu --set full` capture
"""
import collections
import csv
import json
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__grid_size",
        "launch__block_size", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio")
STAGE = {"log_fused_kernel": "log_fused", "log_scan_kernel": "log_scan", "log_hist_kernel": "select_hist",
         "log_rec_kernel": "select_compact", "seg_analyze_kernel": "seg_analyze", "qd_hist_kernel": "qd_hist",
         "qd_collect_kernel": "qd_collect", "qd_rank_kernel": "qd_rank"}


def short(name):
    n = name.replace("void ", "").replace("<unnamed>::", "")
    return n.split("(")[0][:64]


def launches(path, out, cmd):
    rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit()]
    per = collections.defaultdict(list)
    for r in rows:
        if r[12] == "gpu__time_duration.sum":
            per[short(r[4])].append(float(r[14]) / (1e3 if r[13] in ("ns", "nsecond") else 1.0))
    total = sum(sum(v) for v in per.values())
    with open(out, "w") as f:
        f.write(f"# launch list\n\nCommand: `{cmd}`\n(per-launch times are cold-cache and serialised - compare SHARES). "
                f"Raw: {path.split('/')[-1]}\n\n| kernel | launches | avg us | share of device time |\n|---|---:|---:|---:|\n")
        for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"| {k} | {len(v)} | {sum(v) / len(v):.1f} | {100 * sum(v) / total:.1f}% |\n")


def full(path, out, traffic_path, nrows, cmd):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    seen, traffic = set(), {}
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary\n\nCommand: `{cmd}`\n")
        for r in data:
            name = short(r[col["Kernel Name"]])
            if name in seen:
                continue
            seen.add(name)
            f.write(f"\n## {name}\n\n| metric | value |\n|---|---|\n")
            for m in KEEP:
                if m in col:
                    f.write(f"| {m} | {r[col[m]]} {units[col[m]]} |\n")

            def mbytes(metric):
                v, u = float(r[col[metric]].replace(",", "")), units[col[metric]].lower()
                return v * {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1e-6)
            t = mbytes("dram__bytes_read.sum") + mbytes("dram__bytes_write.sum")
            f.write(f"| DRAM traffic (read+write) | {t:.1f} MB |\n")
            for key, stage in STAGE.items():
                if name.startswith(key) and stage not in traffic:
                    traffic[stage] = t * 1e6
    json.dump({"rows": int(nrows), "source": f"{out} (dram__bytes_read.sum + dram__bytes_write.sum per launch)",
               "kernels": traffic}, open(traffic_path, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5], sys.argv[6])
