"""Key metrics + stall ranking of each kernel in an .ncu-rep (via `ncu -i ... --page raw --csv`)."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.per_cycle_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__sass_thread_inst_executed_op_dfma_pred_on.sum', 'sm__sass_thread_inst_executed_op_dmul_pred_on.sum',
        'sm__sass_thread_inst_executed_op_dadd_pred_on.sum', 'sm__cycles_elapsed.avg']
for r in rows[2:]:
    for k in keys:
        for i, h in enumerate(hdr):
            if h == k:
                print(f"{k:70s} {units[i]:14s} {r[i][:90]}")
    st = [(float(r[i].replace(',', '')), h) for i, h in enumerate(hdr)
          if 'smsp__average_warp' in h and 'issue_stalled' in h and h.endswith('.ratio') and 'not_issued' not in h]
    print("warp stalls per issue:", ", ".join(f"{h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} {v:.2f}" for v, h in sorted(st, reverse=True)[:8]))
    print("=====")
