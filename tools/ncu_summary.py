"""Summarise an .ncu-rep (first profiled launch) into the handful of metrics DESIGN.md discusses."""
import csv, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, unit = rows[0], rows[1]
keep = re.compile(r'^(gpu__time_duration.sum|dram__bytes_(read|write).sum|launch__(registers_per_thread|occupancy_limit_\w+|shared_mem_per_block|waves_per_multiprocessor|grid_size|block_size)|sm__throughput.avg.pct_of_peak_sustained_elapsed|sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed|sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)|sm__inst_executed_pipe_(alu|fma|lsu|fp64).avg.pct_of_peak_sustained_active|sm__warps_active.avg.pct_of_peak_sustained_active|smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum|smsp__average_warps_issue_stalled_\w+_per_issue_active.ratio|smsp__warps_eligible.avg.per_cycle_active|smsp__thread_inst_executed_per_inst_executed.ratio|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.avg|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum)$')
for r in rows[2:]:
    print("# kernel:", r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
    for i, h in enumerate(hdr):
        if keep.match(h):
            print('%-95s %-16s %s' % (h, unit[i], r[i]))
