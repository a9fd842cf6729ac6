"""ncu -i X.ncu-rep --page raw --csv | python tools/ncu_extract.py > profiles/<name>.csv — keeps the metrics the summaries quote"""
import csv, sys
KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "launch__grid_size",
        "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg", "sm__inst_executed_pipe_tensor.sum", "launch__occupancy_limit_shared_mem",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio")
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
idx = [i for i, h in enumerate(hdr) if h in KEEP]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [f"launch {k}" for k in range(len(rows) - 2)])
for i in idx:
    w.writerow([hdr[i], units[i]] + [r[i] for r in rows[2:]])
