#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i`, no GPU needed) into the markdown kept under profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--source] > profiles/rNN_name.md
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_drain_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, "raw")
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f"# ncu summary of `{rep}`\n")
    for d in data:
        print(f"## {d[hdr.index('Kernel Name')]}  (launch id {d[hdr.index('ID')]})\n")
        print("| metric | value | unit |\n|---|---:|---|")
        for k in KEYS:
            if k in hdr:
                print(f"| {k} | {d[hdr.index(k)]} | {units[hdr.index(k)]} |")
        print()
    if "--source" in sys.argv:
        src = page(rep, "source")
        while src and (len(src[0]) < 4 or src[0][0].strip() == "Kernel Name"):
            src = src[1:]
        if len(src) > 2:
            h = src[0]
            def col(name):
                for i, x in enumerate(h):
                    if x.strip() == name:
                        return i
                return None
            c_src, c_samp, c_inst = col("Source"), col("# Samples"), col("Instructions Executed")
            if c_src is not None and c_samp is not None:
                body = [r for r in src[1:] if len(r) > max(c_src, c_samp)]
                def num(x):
                    try:
                        return float(x)
                    except ValueError:
                        return 0.0
                tot = sum(num(r[c_samp]) for r in body) or 1.0
                body.sort(key=lambda r: -num(r[c_samp]))
                c_thr = col("Thread Instructions Executed")
                ops, tot_i = {}, 0.0
                for r in body:
                    txt = r[c_src].strip().split()
                    if not txt:
                        continue
                    op = (txt[1] if txt[0].startswith("@") and len(txt) > 1 else txt[0]).split(".")[0]
                    n = num(r[c_inst]) if c_inst is not None else 0.0
                    o = ops.setdefault(op, [0.0, 0.0])
                    o[0] += n
                    o[1] += num(r[c_samp])
                    tot_i += n
                print(f"## SASS opcode mix: {len(body)} static instructions, {tot_i:.0f} warp-instructions executed\n")
                print("| opcode | % of executed warp-instructions | % of stall samples |\n|---|---:|---:|")
                for op, (n, sm) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:25]:
                    print(f"| {op} | {100 * n / (tot_i or 1):.1f} | {100 * sm / tot:.1f} |")
                print()
                print("## hottest lines / instructions by warp-state samples (first launch)\n")
                print("| samples % | instr executed | source |\n|---:|---:|---|")
                for r in body[:40]:
                    print(f"| {100 * num(r[c_samp]) / tot:.1f} | {r[c_inst] if c_inst is not None else ''} | `{r[c_src].strip()[:110]}` |")


if __name__ == "__main__":
    main()
