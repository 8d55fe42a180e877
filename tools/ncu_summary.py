#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): per kernel duration, DRAM bytes, issue utilisation, occupancy,
top stall reasons, and the hottest source lines.   python tools/ncu_summary.py file.ncu-rep [--lines N]"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum"]


def main():
    rep = sys.argv[1]
    n_lines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 25
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[hdr.index("Kernel Name")][:90])
        for k in KEYS:
            if k in hdr:
                print(f"   {k:70s} {r[hdr.index(k)]:>16s} {units[hdr.index(k)]}")
        stalls = [(float(r[i] or 0), hdr[i]) for i in range(len(hdr)) if hdr[i].startswith("smsp__average_warps_issue_stalled") and hdr[i].endswith("_per_issue_active.ratio") or hdr[i].startswith("smsp__average_warp_latency_issue_stalled")]
        for v, name in sorted(stalls, reverse=True)[:8]:
            print(f"   stall {name:75s} {v:.3f}")
    if n_lines:
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(src)))
        hdr = None
        body = []
        fn = ""
        for r in rows:
            if r and r[0] == "Function Name":
                if hdr and body:
                    report(fn, hdr, body, n_lines)
                fn, hdr, body = r[1][:80], None, []
            elif r and r[0] == "Line No":
                hdr = r
            elif hdr and len(r) == len(hdr) and r[0].isdigit():
                body.append(r)
        if hdr and body:
            report(fn, hdr, body, n_lines)


def report(fn, hdr, body, n_lines):
    col = {}
    for i, n in enumerate(hdr):
        col.setdefault(n, i)
    key, ex = "Warp Stall Sampling (All Samples)", "Instructions Executed"
    f = lambda r, c: float(r[col[c]]) if r[col[c]] not in ("", "-") else 0.0
    tot = sum(f(r, key) for r in body) or 1
    tot_ex = sum(f(r, ex) for r in body) or 1
    stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    print(f"-- {fn}: hottest source lines by stall samples (total {tot:.0f}); inst = share of executed warp instructions ({tot_ex:.0f})")
    for r in sorted(body, key=lambda r: -f(r, key))[:n_lines]:
        top = sorted(((f(r, c), c[6:]) for c in stall_cols), reverse=True)[:3]
        tops = " ".join(f"{n}:{v:.0f}" for v, n in top if v > 0)
        print(f"   {100 * f(r, key) / tot:5.1f}%  inst {100 * f(r, ex) / tot_ex:5.1f}%  L{r[0]:>5s}  {r[1].strip()[:100]:100s} [{tops}]")


if __name__ == "__main__":
    main()
