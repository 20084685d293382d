"""Selected metrics of an ncu report as text (one block per captured launch), the format of profiles/r*/..._summary.txt.
usage: python scripts/ncu_summary.py report.ncu-rep "header line" > profiles/rN/name_ncu_full_summary.txt"""
import csv
import subprocess
import sys

KEEP = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate.pct", "launch__block_size", "launch__grid_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "sass__inst_executed_shared_loads", "sass__inst_executed_shared_stores", "sass__inst_executed_global_loads",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active")


def main():
    rep, header = sys.argv[1], sys.argv[2]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    names, units = rows[0], rows[1]
    print("# " + header)
    for n, row in enumerate(rows[2:]):
        kname = row[names.index("Kernel Name")] if "Kernel Name" in names else ""
        print(f"## launch {n} {kname}")
        for k, u, v in sorted(zip(names, units, row)):
            if k in KEEP or (k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")):
                print(k, u, v)


if __name__ == "__main__":
    main()
