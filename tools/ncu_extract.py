"""Summarise an `ncu --set full` report (read here, no GPU needed) into profiles/:
   python tools/ncu_extract.py gpurun_out/v8/prof.ncu-rep r01_v8 [P views H W]
writes profiles/<tag>_ncu_summary.csv (one row per captured launch, the metrics the roofline
argument rests on) and profiles/ncu_traffic.json (DRAM bytes per launch of each kernel, read by
bench.py for roofline.traffic)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum",
    "sm__inst_executed.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
]


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    shape = [int(x) for x in sys.argv[3:7]] if len(sys.argv) >= 7 else [100000, 2, 512, 512]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[0]
    body = [r for r in rows[2:] if len(r) == len(hdr)]   # rows[1] holds the units
    units = dict(zip(hdr, rows[1]))
    name_col = hdr.index("Kernel Name")
    cols = [c for c in KEEP if c in hdr]
    out = os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.csv")
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + cols)
        w.writerow(["(unit)"] + [units[c] for c in cols])
        for r in body:
            w.writerow([r[name_col].split("(")[0]] + [r[hdr.index(c)] for c in cols])
    print("wrote", out, len(body), "launches")

    def scale(col):
        u = units.get(col, "").lower()
        return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)

    def num(x):
        return float(x.replace(",", "")) if x not in ("", "n/a") else 0.0

    stage_of = {"render_backward": "render_backward", "render_forward": "render_forward",
                "preprocess_forward": "preprocess", "preprocess_backward": "preprocess_backward",
                "tile_scan": "tile_scan", "scatter": "scatter", "tile_sort_small": "tile_sort"}
    acc = {}
    for r in body:
        k = r[name_col]
        for pat, st in stage_of.items():
            if pat in k:
                rd = num(r[hdr.index("dram__bytes_read.sum")]) * scale("dram__bytes_read.sum")
                wr = num(r[hdr.index("dram__bytes_write.sum")]) * scale("dram__bytes_write.sum")
                a = acc.setdefault(st, [0.0, 0.0, 0, 0.0, 0.0])
                a[0] += rd
                a[1] += wr
                a[2] += 1
                a[3] += num(r[hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")])
                a[4] += num(r[hdr.index("sm__inst_executed.avg.per_cycle_active")])
                break
    traffic = {st: {"dram_bytes_read": int(a[0] / a[2]), "dram_bytes_write": int(a[1] / a[2]), "launches_averaged": a[2],
                    "issue_active_pct": round(a[3] / a[2], 2), "ipc": round(a[4] / a[2], 3),
                    "P": shape[0], "views": shape[1], "H": shape[2], "W": shape[3], "source": os.path.basename(out)}
               for st, a in acc.items()}
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    json.dump(traffic, open(tp, "w"), indent=1)
    print("wrote", tp, json.dumps(traffic)[:400])


if __name__ == "__main__":
    main()
