"""Multi-GPU host logic: one process per GPU, ``torch.distributed`` for the plumbing.

Two partitions of the evaluator path, matching BASELINE.json's configs:

* population sharding (configs 3, 4): equations are independent units.  Rank r evaluates equations
  r, r + world, ...; the token cache is replicated; NO data-path collective.  Results are merged with
  one gather of O(K) scalars per eval (:func:`gather_population`).
* grid-point sharding (config 5): rank r holds rows [lo, hi) of the time axis of every token
  (:func:`grid_slab`).  Pass A gives partial segment Grams -> ONE all-reduce(sum, float64) -> the
  window solves + LASSO run redundantly and deterministically on every rank -> pass B gives partial
  residual sums -> a second all-reduce of 3 doubles per eval (:func:`evaluate_grid_sharded`).

The collectives run on the stream the kernels run on (NCCL through torch.distributed when the
tensors live on the GPU; the same code runs over gloo with host tensors in the CPU tests).
"""
from __future__ import annotations

import numpy as np


# ------------------------------------------------------------------------------------------
# partitions
# ------------------------------------------------------------------------------------------
def population_slice(n_eq: int, rank: int, world: int):
    """indices of the equations rank `rank` evaluates (round-robin: equal term-count mix per rank,
    since populations are usually sorted by age / fitness, not by size)."""
    if not (0 <= rank < world):
        raise ValueError(f'rank {rank} outside world of {world}')
    return list(range(rank, n_eq, world))


def equation_cost(n_terms: int, n_evals: int) -> float:
    """relative streaming cost of one equation: pass A does T(T+3)/2 Gram updates + ~2T term products per
    point, pass B ~T column products and (evals) x T residual updates at ~1.7x the cost per update
    (DESIGN.md section 6).  Only ratios matter."""
    T = int(n_terms)
    return T * (T + 3) / 2 + 2 * T + 1.7 * T * (int(n_evals) + 1.5)


def population_partition(term_counts, eval_counts, world: int):
    """cost-balanced version of :func:`population_slice`: every rank's equation list (ascending indices) from a
    longest-processing-time greedy on :func:`equation_cost` -- deterministic, identical on every rank.
    Populations of a few dozen equations per rank are otherwise up to ~25 % out of balance (the step time is the
    slowest rank's)."""
    if world <= 0:
        raise ValueError(f'world of {world} ranks')
    if len(term_counts) != len(eval_counts):
        raise ValueError('one eval count per equation expected')
    cost = [equation_cost(t, v) for t, v in zip(term_counts, eval_counts)]
    order = sorted(range(len(cost)), key=lambda i: (-cost[i], i))
    load = [0.0] * world
    parts = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda q: (load[q], len(parts[q]), q))
        parts[r].append(i)
        load[r] += cost[i]
    return [sorted(p) for p in parts]


def grid_slab(shape, rank: int, world: int):
    """(lo, hi) rows of the time axis held by `rank`, the local inner shape and the ``shard``
    argument of :class:`epde_b200.Engine`.  Every rank gets at least 2 rows (engine minimum)."""
    n0 = int(shape[0])
    if not (0 <= rank < world):
        raise ValueError(f'rank {rank} outside world of {world}')
    if n0 < 2 * world:
        raise ValueError(f'time axis of {n0} rows cannot be split over {world} ranks (>= 2 rows each)')
    lo, hi = rank * n0 // world, (rank + 1) * n0 // world
    return (lo, hi), (hi - lo,) + tuple(int(s) for s in shape[1:]), (lo, n0)


# ------------------------------------------------------------------------------------------
# grid-point sharding: the staged path with its two all-reduces
# ------------------------------------------------------------------------------------------
def evaluate_grid_sharded(engine, enc, seg, sums, group=None, want_sw=True, want_l2=True, sol=None):
    """One evaluation of `enc` on an engine that holds a slab of the grid.

    seg, sums: float64 tensors of ``batch.gram_size`` / ``batch.residual_size`` elements on the
    engine's device (scratch; reusable across calls); sol: optional scratch of ``batch.solution_size(want_sw)``
    elements.  Returns the BatchResult (identical on every rank).

    Collectives: all-reduce of the segment tables (pass A), all-reduce of the solution block (every rank solves
    1 / world of the evals -- the window solves and the LASSO do not shrink with the slab, so solving all of them on
    every rank would be the part of the step that does not scale), all-reduce of the residual sums (pass B)."""
    import torch.distributed as dist
    if seg.is_cuda and hasattr(engine, 'set_stream'):
        # the collectives order against torch's CURRENT stream: run the kernels on that stream too, otherwise the
        # all-reduce could start before pass A / pass B have finished (or the solve before the all-reduce)
        import torch
        engine.set_stream(torch.cuda.current_stream(seg.device).cuda_stream)
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    with engine.batch(enc) as b:
        if seg.numel() < b.gram_size or sums.numel() < b.residual_size:
            raise ValueError(f'scratch too small: seg {seg.numel()} < {b.gram_size} or '
                             f'sums {sums.numel()} < {b.residual_size}')
        seg_v, sums_v = seg[:b.gram_size], sums[:b.residual_size]
        if not want_l2:
            b.set_norms(want_l2=False)
        b.gram(seg_v.data_ptr())
        dist.all_reduce(seg_v, op=dist.ReduceOp.SUM, group=group)
        if world > 1 and hasattr(b, 'solve_part'):
            n_sol = b.solution_size(want_sw)
            if sol is None or sol.numel() < n_sol:
                sol = seg.new_empty(n_sol)
            sol_v = sol[:n_sol]
            b.solve_part(seg_v.data_ptr(), rank, world, sol_v.data_ptr(), with_sw=want_sw)
            dist.all_reduce(sol_v, op=dist.ReduceOp.SUM, group=group)
            b.set_solution(sol_v.data_ptr(), with_sw=want_sw)
        else:
            b.solve(seg_v.data_ptr())
        b.residual(sums_v.data_ptr())
        dist.all_reduce(sums_v, op=dist.ReduceOp.SUM, group=group)
        return b.fetch(sums_v.data_ptr(), want_sw=want_sw)


# ------------------------------------------------------------------------------------------
# population sharding: merge of the per-rank results
# ------------------------------------------------------------------------------------------
def local_evals(evals, my_eqs):
    """evals of the global list that belong to this rank, re-indexed to the local equation list.
    Returns (local evals, positions in the global eval list)."""
    where = {e: i for i, e in enumerate(my_eqs)}
    loc, pos = [], []
    for v, (e, t) in enumerate(evals):
        if e in where:
            loc.append((where[e], int(t)))
            pos.append(v)
    return loc, pos


_FIELDS = ('coef', 'intercept', 'fitness_l1', 'fitness_l2', 'stability', 'n_active', 'all_zero', 'iters')


def gather_population(result, positions, n_eval_total, group=None):
    """all-gather the per-rank BatchResult fields into global eval order: dict field -> array.
    (O(K) scalars per eval; sliding-window weights stay on the rank that computed them.)"""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    payload = (list(positions), {k: np.asarray(getattr(result, k)) for k in _FIELDS})
    parts = [None] * world
    dist.all_gather_object(parts, payload, group=group)
    out = {}
    for k in _FIELDS:
        proto = payload[1][k]
        out[k] = np.zeros((n_eval_total,) + proto.shape[1:], dtype=proto.dtype)
    seen = np.zeros(n_eval_total, dtype=bool)
    for pos, fields in parts:
        pos = np.asarray(pos, dtype=np.int64)
        if len(pos) == 0:
            continue
        if seen[pos].any():
            raise RuntimeError('an eval was evaluated by two ranks')
        seen[pos] = True
        for k in _FIELDS:
            out[k][pos] = fields[k][:len(pos)]
    if not seen.all():
        raise RuntimeError(f'{int((~seen).sum())} evals were evaluated by no rank')
    return out
