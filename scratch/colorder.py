"""offline: how many 8x8 column-block pairs does pass A need for the bench population under (a) the shipped column
order (first appearance, equations by size) and (b) a local-search order that clusters co-occurring columns"""
import itertools, random, sys
import numpy as np
from epde_b200 import synthetic as syn

def population(kind='default', pop=128, seed=5):
    if kind == 'default':
        return syn.make_population(7, pop, seed=seed)
    return syn.make_population_wide(7, pop, int(kind[4:]), seed=seed)

def columns(eqs):
    order = sorted(range(len(eqs)), key=lambda e: -len(eqs[e]))     # stable, most terms first
    col_of, cols_eq = {}, {}
    for e in order:
        for term in eqs[e]:
            key = tuple(sorted(term))
            if key not in col_of:
                col_of[key] = len(col_of)
        cols_eq[e] = [col_of[tuple(sorted(t))] for t in eqs[e]]
    return len(col_of), [cols_eq[e] for e in range(len(eqs))]

def block_pairs(perm, ncols, cliques):
    """perm[c] = position of column c; the ones column is position ncols (last)"""
    nb = (ncols + 1 + 7) // 8
    need = np.zeros((nb, nb), dtype=bool)
    ones_blk = ncols // 8
    for cl in cliques:
        blks = sorted({perm[c] // 8 for c in cl} | {ones_blk})
        for i in blks:
            for j in blks:
                if i <= j:
                    need[i, j] = True
    return int(need.sum())

def local_search(ncols, cliques, iters=20000, seed=0):
    rng = random.Random(seed)
    perm = list(range(ncols))
    best = block_pairs(perm, ncols, cliques)
    for it in range(iters):
        a, b = rng.randrange(ncols), rng.randrange(ncols)
        if perm[a] // 8 == perm[b] // 8:
            continue
        perm[a], perm[b] = perm[b], perm[a]
        cur = block_pairs(perm, ncols, cliques)
        if cur <= best:
            best = cur
        else:
            perm[a], perm[b] = perm[b], perm[a]
    return best, perm

for kind in sys.argv[1:] or ['default']:
    eqs = population(kind)
    ncols, cliques = columns(eqs)
    base = block_pairs(list(range(ncols)), ncols, cliques)
    nb = (ncols + 8) // 8
    best, _ = local_search(ncols, cliques)
    print(f'{kind}: {ncols} columns, {nb} blocks ({nb * (nb + 1) // 2} block pairs at most): shipped order {base}, local search {best}')

# ---- cheap heuristics
def rcm_order(ncols, cliques):
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import reverse_cuthill_mckee
    r, c = [], []
    for cl in cliques:
        for a in cl:
            for b in cl:
                if a != b:
                    r.append(a); c.append(b)
    A = coo_matrix((np.ones(len(r)), (r, c)), shape=(ncols, ncols)).tocsr()
    order = reverse_cuthill_mckee(A, symmetric_mode=True)
    perm = [0] * ncols
    for pos, col in enumerate(order):
        perm[col] = pos
    return perm

def greedy_blocks(ncols, cliques):
    """fill blocks of 8 one after the other: seed = unplaced column with most co-occurrences left, then add the unplaced
    column that shares the most equations with the block"""
    co = np.zeros((ncols, ncols))
    for cl in cliques:
        for a in cl:
            for b in cl:
                if a != b:
                    co[a, b] += 1
    unplaced = set(range(ncols))
    perm, pos = [0] * ncols, 0
    while unplaced:
        seed = max(unplaced, key=lambda c: co[c, list(unplaced)].sum())
        blk = [seed]; unplaced.discard(seed)
        while len(blk) < 8 and unplaced:
            nxt = max(unplaced, key=lambda c: co[c, blk].sum())
            blk.append(nxt); unplaced.discard(nxt)
        for c in blk:
            perm[c] = pos; pos += 1
    return perm

for kind in sys.argv[1:] or ['default']:
    eqs = population(kind)
    ncols, cliques = columns(eqs)
    print(kind, 'rcm', block_pairs(rcm_order(ncols, cliques), ncols, cliques), 'greedy blocks', block_pairs(greedy_blocks(ncols, cliques), ncols, cliques))
    for seed in (1, 2, 3):
        for iters in (500, 2000, 5000):
            print('  local search', iters, local_search(ncols, cliques, iters=iters, seed=seed)[0], end=';')
        print()

def local_search_from(perm, ncols, cliques, iters, seed=0):
    rng = random.Random(seed)
    perm = list(perm)
    best = block_pairs(perm, ncols, cliques)
    for it in range(iters):
        a, b = rng.randrange(ncols), rng.randrange(ncols)
        if perm[a] // 8 == perm[b] // 8:
            continue
        perm[a], perm[b] = perm[b], perm[a]
        cur = block_pairs(perm, ncols, cliques)
        if cur <= best:
            best = cur
        else:
            perm[a], perm[b] = perm[b], perm[a]
    return best

print('--- greedy + short local search')
for kind in sys.argv[1:] or ['default']:
    for seed in (5, 6, 7, 8):
        eqs = population(kind, seed=seed)
        ncols, cliques = columns(eqs)
        g = greedy_blocks(ncols, cliques)
        print(kind, 'pop seed', seed, 'cols', ncols, 'shipped', block_pairs(list(range(ncols)), ncols, cliques), 'greedy', block_pairs(g, ncols, cliques),
              '+300', local_search_from(g, ncols, cliques, 300), '+1000', local_search_from(g, ncols, cliques, 1000))
