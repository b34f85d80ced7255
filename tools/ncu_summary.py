#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the text committed under profiles/.

    python tools/ncu_summary.py gpurun_out/X.ncu-rep [--out profiles/X.txt] [--source-top 25]
"""
import argparse
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_fmaheavy.sum",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_uniform.sum",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
    "smsp__cycles_active.avg", "smsp__warp_issue_stalled_barrier_per_warp_active.pct",
    "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_wait_per_warp_active.pct",
    "smsp__warp_issue_stalled_not_selected_per_warp_active.pct", "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct",
    "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_no_instruction_per_warp_active.pct",
    "smsp__warp_issue_stalled_dispatch_stall_per_warp_active.pct", "smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct",
    "smsp__warp_issue_stalled_selected_per_warp_active.pct",
]


def ncu(args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--out")
    ap.add_argument("--source-top", type=int, default=25)
    a = ap.parse_args()
    out = []
    rows = list(csv.reader(io.StringIO(ncu(["-i", a.rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        out.append(f"== kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
        for k in KEYS:
            if k in d:
                out.append(f"{k:78s} {d[k]:>18s} {units[hdr.index(k)]}")
    # per-source-line hot spots (needs -lineinfo and --import-source on)
    src = ncu(["-i", a.rep, "--page", "source", "--csv"])
    try:
        rows = list(csv.reader(io.StringIO(src)))
        h = rows[0]
        col = {n: i for i, n in enumerate(h)}
        samp = next((n for n in h if n.startswith("# Samples") or n == "Warp Stall Sampling (All Samples)" or "Samples" in n), None)
        if samp:
            body = [r for r in rows[1:] if len(r) == len(h)]
            def num(x):
                try:
                    return float(x.replace(",", ""))
                except ValueError:
                    return 0.0
            body.sort(key=lambda r: -num(r[col[samp]]))
            tot = sum(num(r[col[samp]]) for r in body) or 1.0
            out.append(f"== top source lines by '{samp}' (total {tot:.0f})")
            srccol = col.get("Source", 1)
            for r in body[: a.source_top]:
                out.append(f"{100 * num(r[col[samp]]) / tot:6.2f}%  {r[srccol][:150]}")
    except Exception as e:  # keep the raw-metric part even if the source page is unavailable
        out.append(f"(source page unavailable: {e})")
    text = "\n".join(out) + "\n"
    if a.out:
        open(a.out, "w").write(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
