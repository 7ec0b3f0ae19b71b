#!/usr/bin/env python
"""Summarises an .ncu-rep (raw + source pages) for the trace kernel: key counters and instruction share by region."""
import csv, subprocess, sys, io

def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))

def main(rep):
    raw = page(rep, "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "sm__inst_issued.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
            "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__warps_eligible.avg.per_cycle_active",
            "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
            "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
    print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"  {w} [{units[i]}] = {vals[i]}")
    src = page(rep, "source")
    h = src[1]; data = src[2:]
    ia, isrc, iex, ith = h.index("Address"), h.index("Source"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
    tot = sum(int(r[iex]) for r in data)
    # the unrolled sweep body = the longest run of instructions with (nearly) the maximum execution count
    mx = max(int(r[iex]) for r in data)
    hot = [int(r[iex]) for r in data if int(r[iex]) >= 0.9 * mx]
    n_hot = len(hot)
    print(f"  total warp-instructions = {tot:.4e}; hot-loop instructions (exec >= 0.9 max): {n_hot} SASS lines, share = {sum(hot)/tot*100:.2f}% ; block executions = {mx:.4e}")
    ffma = sum(int(r[iex]) for r in data if r[isrc].strip().split()[0] in ("FFMA", "FMUL", "FADD") and int(r[iex]) >= 0.9 * mx)
    print(f"  FP32 instructions inside the hot loop: share of all = {ffma/tot*100:.2f}%")

if __name__ == "__main__":
    main(sys.argv[1])
