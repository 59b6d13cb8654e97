"""Compact view of one `ncu --page raw --csv` export: a few utilisation metrics and the stall reasons per issue.
usage: python profiles/brief_ncu.py full_raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
keys = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.avg',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__grid_size', 'launch__block_size', 'smsp__average_warp_latency_per_inst_issued.ratio']
for r in rows[2:]:
    print('==', r[hdr.index('Kernel Name')])
    for k in keys:
        if k in hdr:
            print(f'   {k} = {r[hdr.index(k)]}')
    st = [(h, r[hdr.index(h)]) for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
    st = sorted(((float(v.replace(',', '')), h) for h, v in st if v not in ('', 'nan', '-nan')), reverse=True)
    print('   stalls per issue:', ', '.join(f"{h.split('stalled_')[1].split('_per_issue')[0]} {v:.2f}" for v, h in st[:8]))
