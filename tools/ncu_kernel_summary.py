#!/usr/bin/env python
"""Summarise one captured launch of a .ncu-rep (ncu --set full) as text:
   python tools/ncu_kernel_summary.py report.ncu-rep [note ...] > profiles/xyz.txt
Reads the raw page through `ncu -i ... --page raw --csv` and keeps the metrics
the roofline discussion uses (duration, DRAM bytes, occupancy, pipe
utilisation, stall sampling)."""
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg",
    "TPC.TriageCompute.sm__pipe_fp64_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warp_latency_per_inst_issued.ratio",
    "smsp__inst_executed.sum", "smsp__cycles_active.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "local_load_bytes", "smsp__inst_executed_op_local_ld.sum",
    "smsp__inst_executed_op_local_st.sum",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for note in sys.argv[2:]:
        print(note)
    for vals in rows[2:]:
        rec = dict(zip(hdr, zip(units, vals)))
        name = rec.get("Kernel Name", ("", "?"))[1]
        print(f"kernel: {name}")
        for k in KEEP:
            if k in rec:
                u, v = rec[k]
                print(f"  {k:95s} {v} {u}")
        stalls = sorted(((float(v or 0), h) for h, (u, v) in rec.items()
                         if h.startswith("smsp__pcsamp_warps_issue_stalled_")
                         and not h.endswith("_not_issued")), reverse=True)
        tot = sum(s for s, _ in stalls) or 1
        print("  warp-state sampling (share of samples):")
        for sv, h in stalls[:7]:
            print(f"    {h.replace('smsp__pcsamp_warps_issue_stalled_', ''):28s} "
                  f"{100 * sv / tot:5.1f} %")


if __name__ == "__main__":
    main()
