"""Summarise an ncu launch list (csv) and a full-set report (.ncu-rep) into a text file.
usage: python profiles/summarize_ncu.py <title> <launches.csv|-> <report.ncu-rep|-> > profiles/<name>.txt"""
import collections
import csv
import subprocess
import sys

title, launches, rep = sys.argv[1:4]
print('#', title)
if launches != '-':
    print('# (a) launch list: ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)')
    rows = [r for r in csv.reader(open(launches)) if len(r) > 10]
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
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f'{n:70s} launches {c:4d} {t / 1e6:10.3f} ms {100 * t / tot:5.1f}%')
if rep != '-':
    print('# (b) ncu --set full --clock-control none --import-source on: selected metrics per captured launch')
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
            'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
            'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
            'sm__warps_active.avg.pct_of_peak_sustained_active',
            'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
            'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__grid_size', 'launch__block_size',
            'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
    for r in rows[2:]:
        print('; '.join(f'{w}={r[hdr.index(w)]} {units[hdr.index(w)]}' for w in want if w in hdr))
