"""Prints the handful of ncu metrics this project tracks from a `--page raw --csv` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__warps_active.avg.per_cycle_active','launch__registers_per_thread','launch__waves_per_multiprocessor','smsp__inst_executed.sum',
 'smsp__issue_active.avg.pct_of_peak_sustained_active','sm__cycles_elapsed.max','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','lts__t_sector_hit_rate.pct','l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
 'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum','smsp__inst_executed_op_local_ld.sum','smsp__inst_executed_op_local_st.sum']
for r in rows[2:]:
    print('====', r[hdr.index('Kernel Name')][:50])
    for k in keys:
        if k in hdr: print("  %-70s %16s %s" % (k, r[hdr.index(k)][:16], units[hdr.index(k)]))
    st = []
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and h.endswith('per_issue_active.ratio') and 'not_issued' not in h:
            try: st.append((float(r[i]), h.split('issue_stalled_')[1].replace('_per_issue_active.ratio', '')))
            except ValueError: pass
    print("  stalls (warps per issue):", ", ".join("%s %.1f" % (n, v) for v, n in sorted(st, reverse=True)[:7]))
