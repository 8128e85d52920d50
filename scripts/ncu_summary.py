#!/usr/bin/env python
"""Print the handful of ncu raw-page metrics that decide a roofline (run where ncu is installed)."""
import csv, subprocess, sys
KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_uniform.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'smsp__cycles_elapsed.avg.per_second',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed_op_shared_st.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for rep in sys.argv[1:]:
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('==', rep)
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f'  {k:75s} {r[i][:60]:>20s} {units[i]}')
        # tensor-ish metrics by substring
        for i, h in enumerate(hdr):
            if ('tensor' in h or 'tmem' in h.lower() or 'utc' in h.lower()) and h not in KEYS and r[i] not in ('', '0', 'n/a') \
                    and 'peak_sustained' not in h.replace('pct_of_peak_sustained', '') and ('.avg' in h or '.sum' in h) and 'ops_path' not in h:
                print(f'  {h:75s} {r[i][:60]:>20s} {units[i]}')
