"""Condenses an `ncu --set full` report (exported with `ncu -i X.ncu-rep --page raw --csv`) to the metrics the
roofline discussion in DESIGN.md uses.  usage: python tools/ncu_summary.py raw.csv > profiles/<name>.txt"""
import csv
import sys

KEYS = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'sm__cycles_elapsed.avg.per_second',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum', 'l1tex__m_l1tex2xbar_write_bytes_mem_global_op_tma_st.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_uniform.sum', 'smsp__inst_executed.sum', 'smsp__inst_executed_op_tma_ld.sum',
        'smsp__inst_executed_op_tma_st.sum', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__sass_inst_executed_op_shared_ld.sum',
        'smsp__sass_inst_executed_op_shared_st.sum']
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(hdr, zip(units, r)))
    for k in KEYS:
        if k in d:
            print('%-90s %-12s %s' % (k, d[k][0], d[k][1]))
    print()
