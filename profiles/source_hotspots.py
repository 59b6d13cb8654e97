"""Aggregate `ncu --page source --csv --print-source cuda,sass` output by CUDA source line.
usage: python profiles/source_hotspots.py src.csv [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
sections, cur, hdr = [], None, None
for r in rows:
    if len(r) >= 2 and r[0] == 'Function Name':
        cur = {'name': r[1], 'rows': []}
        sections.append(cur)
    elif len(r) > 8 and r[0] == 'Line No':
        hdr = r
    elif cur is not None and len(r) > 8:
        cur['rows'].append(r)
i_samp, i_inst = hdr.index('# Samples'), hdr.index('Instructions Executed')
for sec in sections:
    agg = collections.defaultdict(lambda: [0, 0, ''])
    tot_i = tot_s = 0
    for r in sec['rows']:
        try:
            ln, inst, smp = int(r[0]), int(r[i_inst] or 0), int(r[i_samp] or 0)
        except ValueError:
            continue
        agg[ln][0] += inst
        agg[ln][1] += smp
        agg[ln][2] = r[1][:100]
        tot_i += inst
        tot_s += smp
    print('\n==', sec['name'], 'warp-instructions', tot_i, 'stall samples', tot_s)
    for ln, (i, s, src) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
        print(f'{ln:5d} inst {100 * i / max(1, tot_i):5.1f}% samples {100 * s / max(1, tot_s):5.1f}%  {src}')
