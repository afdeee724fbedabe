#!/usr/bin/env python3
"""Turn the raw ncu outputs of profiles/capture.sh into the committed summaries.

  python profiles/summarize.py launches gpurun_out/launches_r01e.csv profiles/launches_r01e_summary.csv [last_step_launches | step]
  python profiles/summarize.py full gpurun_out/full_r01e.ncu-rep [more.ncu-rep ...] profiles/ncu_full_r01e_summary.csv

`launches`: per-kernel totals and shares of the last `last_step_launches` launches of the list (one step), template
arguments kept.  `full`: one row per captured launch with the counters DESIGN.md argues with (ncu -i --page raw --csv).
"""
import csv
import io
import subprocess
import sys

COLUMNS = [  # (summary column, substring of the ncu raw-page metric name)
    ("duration [ms]", "gpu__time_duration.sum"),
    ("regs [register/thread]", "launch__registers_per_thread"),
    ("achieved_occupancy_pct [%]", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("sm_throughput_pct [%]", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("issue_slots_pct [%]", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
    ("fp64_pipe_pct [%]", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
    ("l1tex_throughput_pct [%]", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l1_data_wavefronts_pct [%]", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
    ("l1_hit_pct [%]", "l1tex__t_sector_hit_rate.pct"),
    ("l2_throughput_pct [%]", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l2_hit_pct [%]", "lts__t_sector_hit_rate.pct"),
    ("dram_throughput_pct [%]", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("stall_long_scoreboard [warps/issue]", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    ("stall_math_pipe [warps/issue]", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"),
    ("stall_not_selected [warps/issue]", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"),
    ("alu_pipe_pct [%]", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("adu_pipe_pct [%]", "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active"),
    ("smem_wavefronts_pct [%]", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
    ("dram_read [Gbyte]", "dram__bytes_read.sum"),
    ("dram_write [Gbyte]", "dram__bytes_write.sum"),
]
TO_GBYTE = {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0, "Tbyte": 1e3}
TO_MS = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def launches(src, dst, last="step"):
    rows = []
    with open(src) as fh:
        lines = [ln for ln in fh if ln.startswith('"')]
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", "")) * TO_MS.get(r["Metric Unit"], 1e-6)
        rows.append((r["Kernel Name"], v))
    if last == "step":
        # one step = from one sfcKeysKernel launch to the next; take the last pair at the usual distance
        marks = [i for i, (k, _) in enumerate(rows) if k.startswith("sfcKeys") or "::sfcKeys" in k]
        gaps = [b - a for a, b in zip(marks, marks[1:])]
        usual = max(set(gaps), key=gaps.count)
        a = [m for m, g in zip(marks, gaps) if g == usual][-1]
        rows = rows[a:a + usual]
    else:
        rows = rows[-int(last):]
    tot, cnt = {}, {}
    for k, v in rows:
        tot[k] = tot.get(k, 0.0) + v
        cnt[k] = cnt.get(k, 0) + 1
    total = sum(tot.values())
    with open(dst, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel", "launches", "total_ms", "share_pct"])
        for k in sorted(tot, key=tot.get, reverse=True):
            w.writerow([k, cnt[k], f"{tot[k]:.4f}", f"{100 * tot[k] / total:.2f}"])
        w.writerow(["TOTAL", len(rows), f"{total:.4f}", "100"])
    print(f"{dst}: {len(rows)} launches, {total:.2f} ms")


def full(reps, dst):
    out_rows = []
    for rep in reps:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rd = list(csv.reader(io.StringIO(raw)))
        hdr, units = rd[0], rd[1]
        for r in rd[2:]:
            row = [r[hdr.index("Kernel Name")]]
            for name, key in COLUMNS:
                idx = [i for i, h in enumerate(hdr) if h == key] + [i for i, h in enumerate(hdr) if h.endswith("." + key)]
                idx = [i for i in idx if r[i] not in ("", "n/a")]
                if not idx:
                    row.append("")
                    continue
                i = idx[0]
                v = float(r[i].replace(",", ""))
                if "Gbyte" in name:
                    v *= TO_GBYTE.get(units[i], 1.0)
                if "[ms]" in name:
                    v *= TO_MS.get(units[i], 1.0)
                row.append(f"{v:.6f}")
            out_rows.append(row)
    with open(dst, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [c[0] for c in COLUMNS])
        w.writerows(out_rows)
    print(f"{dst}: {len(out_rows)} launches")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "step")
    else:
        full(sys.argv[2:-1], sys.argv[-1])
