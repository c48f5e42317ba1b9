"""Prints the handful of raw metrics we track from an .ncu-rep (first kernel)."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
h, unit, v = r[0], r[1], r[2]
KEYS = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.per_cycle_active',
        'sm__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_fp16.sum',
        'sm__inst_executed_pipe_lsu.sum', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fmaheavy.sum',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active']
for k in KEYS:
    if k in h:
        i = h.index(k); print(f"{k:70s} {v[i]:>18s} {unit[i]}")
for i, k in enumerate(h):
    if 'warp_issue_stalled' in k and k.endswith('_per_warp_active.pct') and 'not_issued' not in k:
        try:
            if float(v[i]) > 3: print(f"{k:70s} {v[i]:>18s}")
        except ValueError: pass
