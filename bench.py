#!/usr/bin/env python
"""bench.py -- candidate-equation fitness evaluations per second on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload heat512] [--impl reference]

One "step" = one pass of the evaluator hot path over the whole synthetic population: every
admissible (equation, target) pair through term assembly + Gram of the column dictionary (pass A, FP64
tensor cores), sliding-window solves + LASSO, and the residual norms (pass B).  An "eval" is one such
pair.

Workloads (BASELINE.json configs; --workload):
    heat512   config 5: 512^3 (t,x,y), population 128.  DEFAULT.
    heat256   config 4: 256^3, population 256.
    wave1024  config 3: 1024^2, population 64.
N > 1: grid-point sharding along t with NCCL all-reduces of the segment tables and of the residual sums (strong
scaling) for every workload; --sharding population deals equations to ranks instead (no collective).
The JSON line follows the driver contract; `value` is device-resident throughput, `e2e` goes
through the public host API (host encodings in, host results out, copies inside the timed region).
`--impl reference` times the CPU port of the reference's algorithm (oracle/, numpy, all host
cores) on a bounded sample of the same workload: the Python reference itself cannot travel to the
GPU box and cannot allocate its window copies at these sizes.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'fitness_evals_per_sec'
UNIT = 'evals/s'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='heat512')
    ap.add_argument('--pop', type=int, default=0, help='override the population size')
    ap.add_argument('--seed', type=int, default=5)
    ap.add_argument('--cpu-seconds', type=float, default=20.0, help='budget of the cpu_baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--sharding', default='grid', choices=['grid', 'population'],
                    help="N > 1: 'grid' = grid-point sharding along t + two NCCL all-reduces (default for every workload: the "
                         "cost of the column-dictionary passes follows the grid, not the population); 'population' = equations "
                         "dealt to ranks, token cache replicated, no collective")
    ap.add_argument('--population', default='default', choices=['default', 'distinct', 'wide16', 'wide30'],
                    help="'default' = the seeded mix of SURVEY 8d; 'distinct' = no term shared between equations; "
                         "'wide16' / 'wide30' = stress populations of 16 / 30-term equations (pop 32 unless --pop)")
    ap.add_argument('--norms', default='l1', choices=['l1', 'both'],
                    help="residual norms of pass B: 'l1' = the unit of SURVEY 8d (LASSO + L1 fitness + stability, the "
                         "MOEADD path); 'both' also accumulates L2Fitness' sum of squares")
    return ap.parse_args()


def peaks():
    """(hbm GB/s, source) -- measured copy bandwidth of this pool's B200s."""
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:       # noqa: BLE001
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                 '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:       # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:       # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            parts = [p.strip() for p in r.split(',')]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def build_workload(name, pop_override, seed, population='default'):
    from epde_b200 import synthetic as syn
    shape, pop, kind = syn.CONFIGS[name]
    if pop_override:
        pop = pop_override
    n_tokens = len(syn.token_labels(kind))
    if population == 'distinct':
        eqs = syn.make_population_distinct(n_tokens, pop, seed=seed)
    elif population.startswith('wide'):
        eqs = syn.make_population_wide(n_tokens, pop_override or 32, int(population[4:]), seed=seed)
    else:
        eqs = syn.make_population(n_tokens, pop, seed=seed)
    evals = syn.make_evals(eqs)
    return shape, kind, eqs, evals


def algorithmic_bytes(eqs, n_points):
    """pass A reads each distinct token tensor an equation references once:
    8 * N * U_A per equation."""
    from epde_b200 import synthetic as syn
    return sum(8 * n_points * syn.distinct_tokens(eq) for eq in eqs)


def algorithmic_bytes_pass_b(eqs, evals, res, n_points):
    """pass B re-reads the tokens of the terms some eval still needs (targets and
    surviving features): 8 * N * U_B per equation, U_B <= U_A, known once the LASSO has run."""
    used = [set() for _ in eqs]
    for v, (e, t) in enumerate(evals):
        used[e].add(t)
        feat = [i for i in range(len(eqs[e])) if i != t]
        for j, cf in zip(feat, res.coef[v]):
            if cf != 0:
                used[e].add(j)
    total = 0
    for eq, u in zip(eqs, used):
        total += 8 * n_points * len({f[0] for i in u for f in eq[i]})
    return total


def measured_traffic(workload, kernel, world, pop_override):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture of this workload
    (profiles/ncu_traffic.json), or None when no capture matches the configuration."""
    if world != 1 or pop_override:
        return None
    try:
        with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
            return json.load(f).get(f'{workload}:{kernel}')
    except Exception:       # noqa: BLE001
        return None


def single_call_latency(device, calls=50):
    """wall time of ONE small Engine.evaluate call -- what a per-equation operator call (EqRightPartSelector.apply on
    one equation: its D admissible targets as one batch) pays on a config-1 sized grid (61 x 61), host encodings in,
    host results out; median of `calls` after a warm-up."""
    import epde_b200
    from epde_b200 import synthetic as syn
    shape, _, kind = syn.CONFIGS['wave61']
    tokens = syn.make_tokens(shape, kind, seed=1)
    eqs = syn.make_population(len(tokens), 1, seed=1, t_fresh=(6, 6), pruned_fraction=0.0)
    evals = syn.make_evals(eqs)
    enc = epde_b200.encode(eqs, evals)
    with epde_b200.Engine(shape, device=device) as eng:
        for i, t in enumerate(tokens):
            eng.upload(i, t)
        for _ in range(5):
            eng.evaluate(enc)
        ts = []
        for _ in range(calls):
            t0 = time.perf_counter()
            eng.evaluate(enc)
            ts.append(time.perf_counter() - t0)
    return {'single_equation_call_us': float(np.median(ts) * 1e6), 'grid': 'x'.join(str(n) for n in shape),
            'terms': len(eqs[0]), 'evals': len(evals), 'calls': calls,
            'note': 'Engine.evaluate of one 6-term equation (all admissible targets), median wall time per call'}


# ------------------------------------------------------------------------------------------
# CPU side: the oracle port (numpy, all host cores through BLAS) -- ONLY the cpu_baseline leg and
# the --impl reference arm may execute oracle/.
# ------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


_POOL_STATE = {}


def _cpu_worker(e):
    """one equation, all admissible targets, on ONE thread (the pool runs one worker per core)."""
    from oracle import epde_oracle as orc
    from epde_b200 import synthetic as syn
    st = _POOL_STATE
    eq = st['eqs'][e]
    targets = syn.admissible_targets(eq)
    t0 = time.perf_counter()
    fn = orc.evaluate_equation_streamed if st['big'] else orc.evaluate_equation
    rs = fn(st['tokens'], st['shape'], eq, targets=targets)
    dt = time.perf_counter() - t0
    out = [dict(target=int(r.target), coef=np.asarray(r.coef, dtype=np.float64), intercept=float(r.intercept),
                fitness=float(r.fitness), stability=float(r.stability), all_zero=bool(r.all_zero),
                cond=float(getattr(r, 'cond', float('nan')))) for r in rs]
    return e, dt, out


def cpu_eval_equations(tokens, shape, eqs, eq_ids, budget_s, workers):
    """Evaluate equations eq_ids (all admissible targets each) with the oracle: `workers` single-threaded processes
    (forked: the token arrays are shared copy-on-write), waves of `workers` equations until the time budget is
    spent.  Returns (results {eq: [per-target dicts]}, wall seconds, summed worker seconds)."""
    import multiprocessing as mp
    _POOL_STATE.update(tokens=tokens, shape=tuple(shape), eqs=eqs, big=int(np.prod(shape)) > (1 << 22))
    eq_ids = list(eq_ids)
    results, worker_s = {}, 0.0
    t0 = time.perf_counter()
    if workers <= 1:
        for e in eq_ids:
            _, dt, out = _cpu_worker(e)
            results[e] = out; worker_s += dt
            if time.perf_counter() - t0 >= budget_s:
                break
    else:
        ctx = mp.get_context('fork')
        with ctx.Pool(workers, initializer=_limit_threads) as pool:
            for lo in range(0, len(eq_ids), workers):
                for e, dt, out in pool.map(_cpu_worker, eq_ids[lo:lo + workers]):
                    results[e] = out; worker_s += dt
                if time.perf_counter() - t0 >= budget_s:
                    break
    return results, time.perf_counter() - t0, worker_s


def _limit_threads():
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:       # noqa: BLE001
        pass
    os.environ['OMP_NUM_THREADS'] = '1'


def cpu_workers(shape):
    """processes of the parallel CPU baseline: one per core, bounded by host memory (a streamed 512^3 evaluation
    holds about 2 GB of temporaries per worker)."""
    per_worker = 2.5e9 if int(np.prod(shape)) > (1 << 26) else 0.5e9
    try:
        import psutil
        mem = psutil.virtual_memory().available
    except Exception:       # noqa: BLE001
        mem = 32e9
    return int(max(1, min(host_cores(), mem * 0.6 // per_worker, 32)))


def cpu_baseline(tokens, shape, eqs, budget_s, workload):
    """the CPU port of the reference's algorithm on the box's host cores: (i) ALL cores -- one single-threaded worker
    per core, an equation each -- and (ii) the 1-thread figure (evals per worker-second)."""
    from epde_b200 import synthetic as syn
    workers = cpu_workers(shape)
    results, wall, worker_s = cpu_eval_equations(tokens, shape, eqs, range(len(eqs)), budget_s, workers)
    n_ev = sum(len(v) for v in results.values())
    base = {'value': n_ev / wall, 'unit': UNIT, 'cores': workers, 'kind': 'port',
            'single_thread': {'value': n_ev / worker_s, 'unit': UNIT, 'cores': 1},
            'host_cores': host_cores(),
            'sample': f'{len(results)} of {len(eqs)} equations of {workload} ({n_ev} evals, full grid '
                      f'{"x".join(map(str, shape))}) in {wall:.1f} s wall on {workers} single-threaded worker processes '
                      f'(one equation each; {worker_s:.1f} worker-seconds in total); numpy Gram restatement '
                      f'oracle/epde_oracle.py'}
    return base, results


def parity_against(results, gpu, eqs, evals):
    """GPU results vs the oracle results of the cpu_baseline sample, at the FULL size of the workload: identical
    selected term sets, worst relative deviation of coefficients / intercept / L1 fitness / stability over the
    well-posed evals (window condition number <= 1e9, as in the tests)."""
    where = {(e, t): v for v, (e, t) in enumerate(evals)}
    checked = ill = 0
    same_sets, worst = True, 0.0
    for e, outs in results.items():
        F = len(eqs[e]) - 1
        for o in outs:
            v = where.get((e, o['target']))
            if v is None:
                continue
            if o['cond'] > 1e9:
                ill += 1
                continue
            coef = np.asarray(gpu.coef[v, :F])
            if not np.array_equal(coef != 0, o['coef'] != 0):
                same_sets = False
                continue
            scale = max(1e-300, float(np.abs(o['coef']).max()) if F else 1.0)
            errs = [float(np.abs(coef - o['coef']).max() / scale) if F else 0.0,
                    abs(gpu.fitness_l1[v] - o['fitness']) / max(abs(o['fitness']), 1e-300)]
            if not o['all_zero']:
                errs.append(abs(gpu.intercept[v] - o['intercept']) / max(abs(o['intercept']), 1e-9 * scale, 1e-300))
            worst = max(worst, *errs)
            checked += 1
    return {'checked_evals': checked, 'ill_posed_skipped': ill, 'selected_sets_equal': bool(same_sets),
            'max_rel_err': worst, 'tolerance': 1e-6,
            'against': 'oracle/epde_oracle.py (streamed Gram restatement) on the cpu_baseline sample, full grid'}


def run_reference(args):
    """--impl reference: CPU port of the reference's algorithm on the host cores, same data as the GPU arm (the
    counter-based noise is a pure function of seed / token / point index).  Under torchrun only rank 0 works."""
    if int(os.environ.get('RANK', '0')) != 0:
        return
    from epde_b200 import synthetic as syn
    shape, kind, eqs, evals = build_workload(args.workload, args.pop, args.seed, args.population)
    total_budget = max(30.0, 8.0 * args.cpu_seconds)
    per_step = total_budget / max(1, args.steps + args.warmup)
    tokens = syn.make_tokens(shape, kind, seed=args.seed)
    checksum = sum(syn.data_checksum(t) for t in tokens)
    workers = cpu_workers(shape)
    pos = 0

    def step(budget):
        nonlocal pos
        ids = [(pos + i) % len(eqs) for i in range(len(eqs))]
        results, wall, worker_s = cpu_eval_equations(tokens, shape, eqs, ids, budget, workers)
        pos = (pos + len(results)) % len(eqs)
        return sum(len(v) for v in results.values()), wall, len(results), worker_s

    for _ in range(args.warmup):
        step(min(per_step, 5.0))
    tot_ev, tot_dt, tot_eq, tot_ws = 0, 0.0, 0, 0.0
    for _ in range(args.steps):
        n_ev, dt, n_eq, ws = step(per_step)
        tot_ev += n_ev; tot_dt += dt; tot_eq += n_eq; tot_ws += ws
    value = tot_ev / tot_dt
    sample = (f'{tot_eq} equations ({tot_ev} evals) of the {len(eqs)}-equation population over {args.steps} steps on '
              f'{workers} single-threaded worker processes, full grid {"x".join(map(str, shape))}; numpy Gram restatement '
              f'(oracle/epde_oracle.py): the Python reference cannot travel to the GPU box nor allocate its window copies '
              f'at this size')
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * tot_dt / max(1, args.steps),
            'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic',
            'config': dict(workload_config(args, shape, eqs, evals, 1), data_checksum_rank0=float(f'{checksum:.9e}')),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': workers, 'kind': 'port', 'sample': sample,
                             'single_thread': {'value': tot_ev / max(tot_ws, 1e-9), 'unit': UNIT, 'cores': 1},
                             'host_cores': host_cores()},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


def workload_config(args, shape, eqs, evals, world):
    from epde_b200 import synthetic as syn
    T = [len(e) for e in eqs]
    grid_sharded = getattr(args, 'sharding', 'grid') == 'grid' and world > 1
    return {'workload': args.workload, 'grid': 'x'.join(map(str, shape)), 'population': len(eqs),
            'evals_per_step': len(evals), 'terms_per_equation': f'{min(T)}..{max(T)} (mean {np.mean(T):.2f})',
            'distinct_tokens_mean': float(np.mean([syn.distinct_tokens(e) for e in eqs])),
            'sharding': ('grid-points along t + NCCL all-reduce' if grid_sharded else
                         ('population (cost-balanced equation lists), no collective' if world > 1 else 'single GPU')),
            'l2_policy': 'token set larger than L2' if int(np.prod(shape)) * 8 * 3 > 126e6 else
                         'L2 flushed (256 MB write) between timed steps',
            'norms': 'L1 fitness + stability (L2LRFitness, the unit of SURVEY 8d)' if getattr(args, 'norms', 'l1') == 'l1'
                     else 'L1 + L2 fitness + stability',
            'population_kind': getattr(args, 'population', 'default'),
            'seed': args.seed}


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import epde_b200
    from epde_b200 import synthetic as syn

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if world != args.gpus and rank == 0:
        print(f'[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE', file=sys.stderr)
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    shape, kind, eqs, evals = build_workload(args.workload, args.pop, args.seed, args.population)
    grid_sharded = args.sharding == 'grid' and world > 1
    from epde_b200 import sharding
    if grid_sharded:
        (lo, hi), local_shape, shard = sharding.grid_slab(shape, rank, world)
        my_eqs, my_evals = eqs, evals
    else:
        lo, hi, local_shape, shard = 0, shape[0], tuple(shape), None
        if world == 1:
            my_eqs, my_evals = eqs, evals
        else:
            n_ev = [0] * len(eqs)
            for e, _ in evals:
                n_ev[e] += 1
            my_eqs = [eqs[i] for i in sharding.population_partition([len(q) for q in eqs], n_ev, world)[rank]]
            my_evals = syn.make_evals(my_eqs)
    n_local = int(np.prod(local_shape))

    stream = torch.cuda.current_stream()
    eng = epde_b200.Engine(local_shape, device=local_rank, shard=shard, stream=stream.cuda_stream)
    t_up0 = time.perf_counter()
    n_tok = len(syn.token_labels(kind))
    data_checksum = 0.0
    for i in range(n_tok):
        tk = syn.make_tokens(shape, kind, xp=torch, device=dev, seed=args.seed, axis0_range=(lo, hi), only=[i])[i]
        eng.upload_device(i, tk.data_ptr(), tk.numel())
        data_checksum += float(tk[::4099].sum().item())      # sampled: same data on every box <=> same number
        torch.cuda.synchronize()
        del tk
    torch.cuda.empty_cache()
    t_gen = time.perf_counter() - t_up0

    enc = epde_b200.encode(my_eqs, my_evals)
    flush = None
    if n_local * 8 * 3 <= 126e6:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident step: encoding already on the device, results stay on the device
    batch = eng.batch(enc)
    want_l2 = args.norms == 'both'
    batch.set_norms(want_l2)
    seg = torch.zeros(batch.gram_size, dtype=torch.float64, device=dev)
    sums = torch.zeros(batch.residual_size, dtype=torch.float64, device=dev)

    sol = torch.zeros(batch.solution_size(False), dtype=torch.float64, device=dev) if grid_sharded else None

    def device_step():
        batch.gram(seg.data_ptr())
        if grid_sharded:
            dist.all_reduce(seg)
            # every rank solves 1 / world of the evals; the solution block is merged by a sum
            batch.solve_part(seg.data_ptr(), rank, world, sol.data_ptr())
            dist.all_reduce(sol)
            batch.set_solution(sol.data_ptr())
        else:
            batch.solve(seg.data_ptr())
        batch.residual(sums.data_ptr())
        if grid_sharded:
            dist.all_reduce(sums)

    def timed(step_fn, k):
        """sum of per-step device times (CUDA events on the launching stream), L2 flushed between
        steps when the token set fits L2; barrier + synchronize on both sides."""
        pairs = []
        barrier()
        for _ in range(k):
            if flush is not None:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            step_fn()
            b.record(stream)
            pairs.append((a, b))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in pairs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    info = batch.info()
    for _ in range(args.warmup):
        device_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- the timed region proper: no per-kernel event bracketing
    l0 = eng.launch_count
    ms_total = timed(device_step, args.steps)
    launches = eng.launch_count - l0
    # ---- kernel classes: a second, shorter run with every launch bracketed by events
    k_steps = min(args.steps, 3)
    eng.set_timing(True)
    timed(device_step, k_steps)
    ktimes = eng.kernel_times()
    eng.set_timing(False)
    ktimes = {k: (v[0] / k_steps, v[1] // k_steps) for k, v in ktimes.items()}       # per step
    # per-rank sum of kernel time per step (the step time is the slowest rank's)
    rank_ms = torch.tensor([sum(v[0] for v in ktimes.values())], dtype=torch.float64, device=dev)
    if world > 1:
        gathered = [torch.zeros_like(rank_ms) for _ in range(world)]
        dist.all_gather(gathered, rank_ms)
        rank_ms = torch.cat(gathered)
    rank_ms = [round(float(x), 3) for x in rank_ms.tolist()]
    dev_res = batch.fetch(sums.data_ptr(), want_sw=False)
    batch.close()
    total_evals = len(evals)
    value = total_evals * args.steps / (ms_total * 1e-3)

    # ---- end to end through the host API: host encodings in, host results out (sliding-window weights included:
    #      LASSOSparsity.apply always leaves them on the equation, sparsity.py:210), every step
    def e2e_step():
        if grid_sharded:
            return sharding.evaluate_grid_sharded(eng, enc, seg, sums, want_sw=True, want_l2=want_l2)
        return eng.evaluate(enc, want_sw=True, want_l2=want_l2)

    def timed_wall(step_fn, k):
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            if flush is not None:
                flush.fill_(1)
            res = step_fn()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), res

    e2e_step()
    e2e_s, last = timed_wall(e2e_step, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    h2d = sum(a.nbytes for a in (enc.term_off, enc.fac_off, enc.fac_token, enc.fac_power, enc.fac_kind,
                                 enc.fac_param, enc.eval_eq, enc.eval_target))
    d2h = sum(getattr(last, k).nbytes for k in ('coef', 'intercept', 'fitness_l1', 'fitness_l2', 'stability',
                                                'sw_weights', 'n_active', 'all_zero', 'iters'))
    e2e_value = total_evals * args.steps / e2e_s

    if rank == 0:
        hbm_peak, hbm_src = peaks()
        dfma_tf, dmma_tf = epde_b200.Engine.fp64_peak(local_rank, 1.0)
        nnz = int((np.asarray(dev_res.coef) != 0).sum())
        n_my_evals = len(my_evals)
        # ---- FP64 work per step.  Pass A: `block_pairs` DMMA.8x8x4 (512 flop each) per 4 points = executed; the pair
        #      entries some equation needs (one FMA per point each) = useful.  Pass B: per eval and point one subtract, one
        #      FMA per surviving coefficient, one |.|-accumulate.
        flops = {'gram': {'executed': 128.0 * info['block_pairs'] * n_local, 'useful': 2.0 * info['entries'] * n_local},
                 'residual': {'executed': (2.0 * n_my_evals + 2.0 * nnz) * n_local,
                              'useful': (2.0 * n_my_evals + 2.0 * nnz) * n_local}}
        alg = {'gram': algorithmic_bytes(my_eqs, n_local),
               'residual': algorithmic_bytes_pass_b(my_eqs, my_evals, dev_res, n_local)}
        names = {'gram': 'gram_dmma_kernel (pass A: column dictionary + DMMA.8x8x4 Gram, TMA-staged token tiles)',
                 'residual': 'residual_cols_kernel (pass B: column dictionary + sparse residual sums)'}
        peak_of = {'gram': ('tensor', dmma_tf, 'measured in this run: epde_measure_fp64_peak, DMMA.8x8x4 chains (FP64 tensor pipe)'),
                   'residual': ('fp64', dfma_tf, 'measured in this run: epde_measure_fp64_peak, DFMA chains (FP64 pipe)')}

        def roof(kind):
            ms, n = ktimes[kind]
            bound, peak, src = peak_of[kind]
            ach = flops[kind]['executed'] / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
            useful = flops[kind]['useful'] / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
            traffic = measured_traffic(args.workload, kind, world, args.pop or args.population != 'default' or args.norms != 'l1')
            hbm = {'peak_gbs': hbm_peak, 'peak_source': hbm_src,
                   'algorithmic_bytes_per_step': int(alg[kind]),
                   'algorithmic_gbs': alg[kind] / (ms * 1e-3) / 1e9 if ms > 0 else 0.0,
                   'dram_bytes_measured_per_launch': traffic,
                   'dram_gbs_measured': (traffic / (ms * 1e-3) / 1e9) if traffic and ms > 0 else None,
                   'note': 'SURVEY 8(d) algorithmic bytes (8*N*U per equation) are an accounting figure: the whole launch '
                           'group shares ONE sweep over the token tiles, so the DRAM traffic is the token set read once '
                           '(+ the cell table) and HBM is far from binding'}
            if hbm['dram_gbs_measured']:
                hbm['dram_frac'] = hbm['dram_gbs_measured'] / hbm_peak
            out = {'bound': bound, 'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak if peak else None,
                   'traffic': traffic, 'kernel': names[kind], 'peak_source': src, 'launches_per_step': int(n),
                   'kernel_ms_per_step': ms, 'flop_per_step_executed': flops[kind]['executed'],
                   'flop_per_step_useful': flops[kind]['useful'], 'useful_tflops': useful,
                   'useful_frac': useful / peak if peak else None, 'hbm': hbm}
            # pipe utilisations of the committed ncu --set full capture of this configuration (NOT measured in this
            # run): which on-chip resource binds the kernel
            ncu = measured_traffic(args.workload, kind + ':ncu', world,
                                   args.pop or args.population != 'default' or args.norms != 'l1')
            if ncu:
                out['ncu_capture'] = ncu
            return out

        dominant = 'residual' if ktimes['residual'][0] >= ktimes['gram'][0] else 'gram'
        other = 'gram' if dominant == 'residual' else 'residual'
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_total / args.steps, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': dict(workload_config(args, shape, eqs, evals, world),
                           data_checksum_rank0=float(f'{data_checksum:.9e}'),
                           nonzero_coefficients_rank0=nnz,
                           column_dictionary_rank0=info),
            'clocks': clocks,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
                    'note': 'Engine.evaluate (C ABI epde_evaluate_batch): host CSR encodings in, host results out incl. the '
                            'sliding-window weights; token tensors are cache-resident in HBM as in the reference (filled once '
                            'per dataset)',
                    'token_cache_fill_s': t_gen},
            'gpu_launches': int(launches),
            'roofline': roof(dominant),
            'roofline_other': roof(other),
            'fp64_peaks_tflops': {'dfma': dfma_tf, 'dmma': dmma_tf},
            'kernel_ms_per_step': {k: v[0] for k, v in ktimes.items()},
            'kernel_ms_per_step_by_rank': rank_ms,
        }
        if world == 1:
            line['latency'] = single_call_latency(local_rank)
        if world == 1 and not args.no_cpu_baseline:
            host_tokens = [eng.download(i) for i in range(n_tok)]
            # the host upload path an EPDE run would use (epde_cache_upload: pageable host memory -> HBM), timed once
            t0 = time.perf_counter()
            for i in range(n_tok):
                eng.upload(i, host_tokens[i])
            line['e2e']['token_upload_host_s'] = time.perf_counter() - t0
            line['e2e']['token_upload_host_bytes'] = int(sum(t.nbytes for t in host_tokens))
            base, results = cpu_baseline(host_tokens, shape, eqs, args.cpu_seconds, args.workload)
            line['cpu_baseline'] = base
            line['parity'] = parity_against(results, dev_res, eqs, evals)
            del host_tokens
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
