#!/usr/bin/env python
"""Writes/updates the entry of one workload in profiles/traffic.json from an ncu --set full report.
usage: traffic_entry.py <workload> <report.ncu-rep> <summary file under profiles/> "<how>" """
import csv, io, json, os, subprocess, sys

def main(workload, rep, source, how):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, vals = rows[0], rows[2]
    g = lambda k: float(vals[hdr.index(k)].replace(",", ""))
    unit = lambda k: rows[1][hdr.index(k)]
    def bytes_of(k):
        v, u = g(k), unit(k).lower()
        return int(v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u])
    rd, wr = bytes_of("dram__bytes_read.sum"), bytes_of("dram__bytes_write.sum")
    ms = g("gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(unit("gpu__time_duration.sum"), 1)
    e = {
        "kernel": vals[hdr.index("Kernel Name")], "kernel_ms_under_ncu": round(ms, 3),
        "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": rd + wr,
        "issue_slots_busy_pct": round(g("sm__inst_issued.avg.pct_of_peak_sustained_active"), 1),
        "fma_pipe_pct": round(g("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"), 1),
        "fma_pipe_cycles_active_pct": round(g("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed"), 1),
        "alu_pipe_pct": round(g("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"), 1),
        "lsu_pipe_pct": round(g("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"), 1),
        "threads_per_instruction": round(g("smsp__thread_inst_executed_per_inst_executed.ratio"), 1),
        "warp_instructions": int(g("smsp__inst_executed.sum")),
        "dispatch_stalls_per_issue": round(g("smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio"), 3),
        "achieved_occupancy_pct": round(g("sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
        "registers_per_thread": int(g("launch__registers_per_thread")),
        "how": how, "source": source,
    }
    for k in ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",):
        if k in hdr:
            e["fp64_pipe_pct"] = round(g(k), 1)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "traffic.json")
    t = json.load(open(path))
    t[workload] = e
    json.dump(t, open(path, "w"), indent=1)
    print(workload, json.dumps(e)[:300])

if __name__ == "__main__":
    main(*sys.argv[1:5])
