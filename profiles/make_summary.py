"""Turn the small files profiles/capture.sh leaves under gpurun_out/ into the committed summary of a kernel version.
usage: python profiles/make_summary.py <tag> "<title>"   ->  profiles/r01_<tag>_ncu_summary.txt, ncu_traffic.json"""
import collections
import csv
import json
import sys

tag, title = sys.argv[1], sys.argv[2]
out = ['# ' + title,
       '# (a) launch list: ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --no-cpu-baseline --steps 2 '
       '--warmup 1 (cold-cache, serialised: compare SHARES)']
rows = [r for r in csv.reader(open(f'gpurun_out/launches_heat512_{tag}.csv')) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(',', ''))
    except ValueError:
        continue
    n = r[ki].split('(')[0][:70]
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:12]:
    out.append(f'{n:70s} launches {c:4d} {t / 1e6:10.3f} ms {100 * t / tot:5.1f}%')
out.append('# (b) ncu --set full --clock-control none --import-source on (-k gram_tile|residual_tile -c 2, --steps 1 --warmup 1)')
rows = list(csv.reader(open(f'gpurun_out/full_raw_{tag}.csv')))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__sass_inst_executed_op_shared_ld.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio']
mult = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}
traffic = {}
for r in rows[2:]:
    out.append('; '.join(f'{w}={r[hdr.index(w)]} {units[hdr.index(w)]}'.strip() for w in want if w in hdr))
    name = r[hdr.index('Kernel Name')]
    try:
        rd, wr = float(r[hdr.index('dram__bytes_read.sum')]), float(r[hdr.index('dram__bytes_write.sum')])
        if rd == rd:
            traffic['gram' if 'gram' in name else 'residual'] = int(rd * mult[units[hdr.index('dram__bytes_read.sum')]] +
                                                                    wr * mult[units[hdr.index('dram__bytes_write.sum')]])
    except (ValueError, KeyError):
        pass
    samp = {h.replace('smsp__pcsamp_warps_issue_stalled_', ''): float(v) for h, v in zip(hdr, r)
            if h.startswith('smsp__pcsamp_warps_issue_stalled') and not h.endswith('not_issued') and v not in ('', 'nan')}
    st = sum(samp.values()) or 1.0
    out.append('    stall samples: ' + ', '.join(f'{k} {100 * v / st:.1f}%' for k, v in sorted(samp.items(), key=lambda x: -x[1])[:8]))
out.append('# (c) ncu --metrics (single pass) for residual_tile_kernel, same command (the full set reports nan for this launch)')
rd = wr = None
for r in csv.reader(open(f'gpurun_out/resid_metrics_{tag}.csv')):
    if len(r) > 14 and r[0] != 'ID':
        out.append(f'{r[4]} grid {r[8]} block {r[7]}: {r[12]} = {r[14]} {r[13]}')
        if r[12] == 'dram__bytes_read.sum':
            rd = int(r[14])
        if r[12] == 'dram__bytes_write.sum':
            wr = int(r[14])
if rd is not None:
    traffic['residual'] = rd + wr
open(f'profiles/r01_{tag}_ncu_summary.txt', 'w').write('\n'.join(out) + '\n')
json.dump({'heat512:gram': traffic.get('gram'), 'heat512:residual': traffic.get('residual'),
           '_source': f'profiles/r01_{tag}_ncu_summary.txt (ncu, heat512 population 128, DRAM read + write bytes per launch)'},
          open('profiles/ncu_traffic.json', 'w'), indent=1)
print('\n'.join(out))
