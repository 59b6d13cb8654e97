"""SASS evidence for the shipped library: opcode histogram per kernel and the first occurrence of every instruction
that carries the TMA / mbarrier data movement (UTMALDG, UBLKCP, SYNCS.*) and the FP64 tensor-core MMA (DMMA.8x8x4).
usage: cuobjdump -sass epde_b200/libepde_b200.so > /tmp/sass.txt; python profiles/sass_extract.py /tmp/sass.txt > profiles/r02_sass_extract.txt"""
import collections
import re
import sys

WANT = ('gram_dmma_kernel', 'residual_cols_kernel', 'lasso_solve_kernel', 'marginal_slab_kernel', 'poly_derivative_kernel',
        'fd_gradient_kernel')
MARK = re.compile(r'(DMMA|UTMALDG|UBLKCP|SYNCS|UTMAPF|BAR\.|UTMACMDFLUSH)')
SEL = re.compile(r'(DMMA|UTMALDG|UBLKCP|SYNCS|DFMA|DMUL|DADD|LDS|STS|LDG|STG|BAR|LDGSTS|MUFU|SHFL)')
INSTR = re.compile(r'\s*/\*([0-9a-f]{4,6})\*/\s+(@!?U?P\d+\s+)?([A-Za-z0-9_.]+)')

txt = open(sys.argv[1]).read()
print('# cuobjdump -sass epde_b200/libepde_b200.so (sm_100a): opcode histogram per kernel + first occurrence of the')
print('# instructions that carry the TMA / mbarrier data movement and the FP64 tensor-core MMA.')
for part in re.split(r'(?m)^\s*Function : ', txt)[1:]:
    name = part.split('\n', 1)[0].strip()
    if not any(w in name for w in WANT):
        continue
    ops, marked = collections.Counter(), []
    for ln in part.split('\n'):
        m = INSTR.match(ln)
        if not m:
            continue
        ops[m.group(3)] += 1
        if MARK.match(m.group(3)):
            marked.append((m.group(3), ln.strip()))
    fam = collections.Counter()
    for op, c in ops.items():
        fam[op.split('.')[0]] += c
    print(f'\n== {name}: {sum(ops.values())} SASS instructions')
    print('   opcode families: ' + ', '.join(f'{k} {v}' for k, v in fam.most_common(24)))
    print('   selected opcodes: ' + ', '.join(f'{k} {ops[k]}' for k in sorted(ops) if SEL.match(k)))
    seen = set()
    for op, ln in marked:
        if op not in seen:
            seen.add(op)
            print('   ' + ln)
