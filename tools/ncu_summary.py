"""Summarise an ncu report (.ncu-rep) into the text form kept under profiles/:
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep "command line that was profiled" > profiles/rNN_xxx.txt"""
import csv
import io
import subprocess
import sys

KEEP = """Kernel Name|Block Size|Grid Size|gpu__time_duration.sum|launch__registers_per_thread|launch__shared_mem_per_block_static
launch__shared_mem_per_block_dynamic|launch__occupancy_limit_registers|launch__occupancy_limit_shared_mem|launch__occupancy_limit_warps
sm__warps_active.avg.per_cycle_active|sm__warps_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active|sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active|sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active|sm__issue_active.avg.pct_of_peak_sustained_elapsed
smsp__issue_active.avg.per_cycle_active|smsp__inst_executed.sum|smsp__thread_inst_executed_per_inst_executed.ratio
sm__throughput.avg.pct_of_peak_sustained_elapsed|dram__bytes_read.sum|dram__bytes_write.sum|dram__bytes_read.sum.per_second
gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|l1tex__t_sector_hit_rate.pct|lts__t_sector_hit_rate.pct|lts__t_sectors.sum
l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio|smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio|smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio""".replace("\n", "|").split("|")

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
print(f"# ncu --set full --clock-control none --import-source on   (report: {rep})")
print(f"# command: {sys.argv[2] if len(sys.argv) > 2 else '?'}")
print(f"# read with: ncu -i {rep} --page raw --csv")
for li, r in enumerate(rows[2:]):
    print(f"--- launch ID {li}")
    d = dict(zip(hdr, zip(units, r)))
    for k in KEEP:
        if k in d:
            print(f"{k:95s} {d[k][0]:16s} {d[k][1]}")
