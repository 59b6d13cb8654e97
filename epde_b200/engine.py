"""Host side of the B200 fitness-evaluation engine: HBM token cache + batched evaluation.

Mirrors, for the evaluator path only:
  * ``Cache`` of epde/cache/cache.py:141-524           -> :class:`Engine` token table
  * ``Equation.evaluate`` main_structures.py:588-641,
    ``calculate_weights`` supplementary.py:394-471,
    ``PhysicsInformedLasso.fit`` sparsity.py:42-150,
    ``L2LRFitness.apply`` / ``L2Fitness.apply`` fitness.py:27-183 -> :meth:`Engine.evaluate`

Everything numeric happens in libepde_b200.so (hand-written sm_100a CUDA); this module only packs
the equation encoding into CSR arrays and unpacks results.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib

MAX_TERMS = 32         # EPDE_MAX_TERMS
MAX_FACTORS = 4        # EPDE_MAX_FACTORS
KS = MAX_TERMS + 1     # stride of a sliding-window weight row

KIND_POWER, KIND_SIN, KIND_COS, KIND_INVERSE = 0, 1, 2, 3


def _p(arr, typ):
    return None if arr is None else arr.ctypes.data_as(typ)


@dataclass
class Encoding:
    """CSR encoding of a batch (include/epde_b200.h 'Equation encoding')."""
    n_eq: int
    term_off: np.ndarray
    fac_off: np.ndarray
    fac_token: np.ndarray
    fac_power: np.ndarray
    fac_kind: np.ndarray
    fac_param: np.ndarray
    eval_eq: np.ndarray
    eval_target: np.ndarray

    @property
    def n_eval(self):
        return int(self.eval_eq.shape[0])

    def terms_of(self, e):
        return int(self.term_off[e + 1] - self.term_off[e])

    def c_args(self):
        return (self.n_eq, _p(self.term_off, _lib.c_i32p), _p(self.fac_off, _lib.c_i32p),
                _p(self.fac_token, _lib.c_i32p), _p(self.fac_power, _lib.c_f64p),
                _p(self.fac_kind, _lib.c_i32p), _p(self.fac_param, _lib.c_f64p),
                self.n_eval, _p(self.eval_eq, _lib.c_i32p), _p(self.eval_target, _lib.c_i32p))


def encode(equations, evals=None) -> Encoding:
    """equations: [[ [(token_id, power[, kind[, param]]), ...] per term ] per equation];
    evals: [(equation index, target term index)], default = every term of every equation."""
    term_off, fac_off = [0], [0]
    tok, pw, kd, pr = [], [], [], []
    for eq in equations:
        for term in eq:
            for fac in term:
                fac = tuple(fac)
                tok.append(int(fac[0]))
                pw.append(float(fac[1]))
                kd.append(int(fac[2]) if len(fac) > 2 else 0)
                pr.append(float(fac[3]) if len(fac) > 3 else 0.0)
            fac_off.append(len(tok))
        term_off.append(len(fac_off) - 1)
    if evals is None:
        evals = [(e, t) for e, eq in enumerate(equations) for t in range(len(eq))]
    ev = np.asarray(evals, dtype=np.int32).reshape(-1, 2)
    return Encoding(len(equations), np.asarray(term_off, np.int32), np.asarray(fac_off, np.int32),
                    np.asarray(tok, np.int32), np.asarray(pw, np.float64), np.asarray(kd, np.int32),
                    np.asarray(pr, np.float64), np.ascontiguousarray(ev[:, 0]),
                    np.ascontiguousarray(ev[:, 1]))


@dataclass
class BatchResult:
    """Per-eval outputs (include/epde_b200.h epde_evaluate_batch)."""
    coef: np.ndarray          # (n_eval, MAX_TERMS) weights_internal in feature order, zero padded
    intercept: np.ndarray     # (n_eval,)
    fitness_l1: np.ndarray    # (n_eval,)  L2LRFitness value
    fitness_l2: np.ndarray    # (n_eval,)  L2Fitness value (before the penalty division)
    stability: np.ndarray     # (n_eval,)
    sw_weights: np.ndarray | None   # (n_eval, W, KS)
    n_active: np.ndarray
    all_zero: np.ndarray
    iters: np.ndarray

    def cached_weights(self, v):
        """cached_weights_ of eval v: (W, n_active + 1)  (sparsity.py:48,88)."""
        return self.sw_weights[v, :, :int(self.n_active[v]) + 1]


def _alloc_out(n_eval, W, want_sw):
    ne = max(1, n_eval)
    return dict(coef=np.zeros((ne, MAX_TERMS)), intercept=np.zeros(ne), fitness_l1=np.zeros(ne),
                fitness_l2=np.zeros(ne), stability=np.zeros(ne),
                sw_weights=np.zeros((ne, W, KS)) if want_sw else None,
                n_active=np.zeros(ne, np.int32), all_zero=np.zeros(ne, np.int32),
                iters=np.zeros(ne, np.int32))


def _out_args(o):
    return (_p(o['coef'], _lib.c_f64p), _p(o['intercept'], _lib.c_f64p), _p(o['fitness_l1'], _lib.c_f64p),
            _p(o['fitness_l2'], _lib.c_f64p), _p(o['stability'], _lib.c_f64p),
            _p(o['sw_weights'], _lib.c_f64p), _p(o['n_active'], _lib.c_i32p),
            _p(o['all_zero'], _lib.c_i32p), _p(o['iters'], _lib.c_i32p))


def _trim(o, n_eval):
    return BatchResult(**{k: (None if v is None else v[:n_eval]) for k, v in o.items()})


class Engine:
    """One evaluator context = one GPU, one inner grid shape (replaces the tensor cache of
    epde/cache/cache.py for this path; created where globals.init_caches is, globals.py:26-42).

    shard=(axis0_offset, axis0_global): this engine holds a contiguous block of the time axis
    (grid-point sharding)."""

    def __init__(self, inner_shape, device=0, shard=None, stream=None):
        self._lib = _lib.load()
        self.inner_shape = tuple(int(s) for s in inner_shape)
        self.n_points = int(np.prod(self.inner_shape))
        self.device = int(device)
        shape = (C.c_int64 * len(self.inner_shape))(*self.inner_shape)
        h = C.c_void_p()
        _lib.check(self._lib.epde_ctx_create(self.device, len(self.inner_shape), shape, C.byref(h)))
        self._h = h
        if shard is not None:
            _lib.check(self._lib.epde_ctx_set_shard(self._h, int(shard[0]), int(shard[1])))
        if stream is not None:
            self.set_stream(stream)

    # ---- lifetime
    def close(self):
        if getattr(self, '_h', None) is not None and self._h:
            self._lib.epde_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:       # noqa: BLE001
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def synchronize(self):
        _lib.check(self._lib.epde_ctx_synchronize(self._h))

    def set_stream(self, cuda_stream):
        """cuda_stream: integer cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""
        _lib.check(self._lib.epde_ctx_set_stream(self._h, C.c_void_p(int(cuda_stream))))

    # ---- token cache (cache.py:333-490)
    def upload(self, token_id, values):
        a = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
        _lib.check(self._lib.epde_cache_upload(self._h, int(token_id), _p(a, _lib.c_f64p), a.size))

    def upload_device(self, token_id, dev_ptr, n):
        _lib.check(self._lib.epde_cache_upload_device(self._h, int(token_id), C.c_void_p(int(dev_ptr)), int(n)))

    def download(self, token_id):
        out = np.empty(self.n_points)
        _lib.check(self._lib.epde_cache_download(self._h, int(token_id), _p(out, _lib.c_f64p), out.size))
        return out

    def __contains__(self, token_id):
        r = C.c_int(0)
        _lib.check(self._lib.epde_cache_contains(self._h, int(token_id), C.byref(r)))
        return bool(r.value)

    def delete(self, token_id):
        _lib.check(self._lib.epde_cache_delete(self._h, int(token_id)))

    @property
    def consumed_memory(self):
        r = C.c_int64(0)
        _lib.check(self._lib.epde_cache_bytes(self._h, C.byref(r)))
        return int(r.value)

    def set_sample_weights(self, g):
        a = np.ascontiguousarray(g, dtype=np.float64).reshape(-1)
        _lib.check(self._lib.epde_ctx_set_sample_weights(self._h, _p(a, _lib.c_f64p), a.size))

    # ---- geometry
    @property
    def n_windows(self):
        r = C.c_int(0)
        _lib.check(self._lib.epde_ctx_num_windows(self._h, C.byref(r)))
        return int(r.value)

    @property
    def n_segments(self):
        r = C.c_int(0)
        _lib.check(self._lib.epde_ctx_num_segments(self._h, C.byref(r)))
        return int(r.value)

    @property
    def launch_count(self):
        r = C.c_int64(0)
        _lib.check(self._lib.epde_ctx_launch_count(self._h, C.byref(r)))
        return int(r.value)

    @staticmethod
    def fp64_peak(device=0, seconds=1.0):
        """(DFMA TFLOP/s, DMMA TFLOP/s) sustained on `device` (include/epde_b200.h epde_measure_fp64_peak)."""
        a, b = C.c_double(0), C.c_double(0)
        lib = _lib.load()
        _lib.check(lib.epde_measure_fp64_peak(int(device), float(seconds), C.byref(a), C.byref(b)))
        return a.value, b.value

    # ---- measurement aid (include/epde_b200.h epde_ctx_set_timing)
    KERNEL_CLASSES = ('gram', 'marginal', 'solve', 'residual')

    def set_timing(self, enabled=True):
        _lib.check(self._lib.epde_ctx_set_timing(self._h, int(bool(enabled))))

    def kernel_times(self):
        """{class: (summed device ms, launches)} since timing was enabled (synchronises)."""
        out = {}
        for k, name in enumerate(self.KERNEL_CLASSES):
            ms, n = C.c_double(0), C.c_int64(0)
            _lib.check(self._lib.epde_ctx_kernel_time(self._h, k, C.byref(ms), C.byref(n)))
            out[name] = (ms.value, int(n.value))
        return out

    # ---- evaluation
    def evaluate(self, enc: Encoding, want_sw=True, want_l2=True) -> BatchResult:
        """Whole path, host buffers in/out, synchronous.  want_l2=False: pass B accumulates the L1 sums
        only (what L2LRFitness needs, fitness.py:142); ``fitness_l2`` is then NaN."""
        if not want_l2:
            with self.batch(enc) as b:
                b.set_norms(want_l2=False)
                b.gram(); b.solve(); b.residual()
                return b.fetch(want_sw=want_sw)
        o = _alloc_out(enc.n_eval, self.n_windows, want_sw)
        _lib.check(self._lib.epde_evaluate_batch(self._h, *enc.c_args(), *_out_args(o)))
        return _trim(o, enc.n_eval)

    def batch(self, enc: Encoding) -> 'Batch':
        return Batch(self, enc)


class Batch:
    """Staged evaluation (epde_batch_*): gram -> [all-reduce] -> solve -> residual -> [all-reduce]
    -> fetch.  Device buffers are passed as integer device pointers (torch ``data_ptr()``)."""

    def __init__(self, engine: Engine, enc: Encoding):
        self.engine = engine
        self.enc = enc          # keeps the host arrays alive
        self._lib = engine._lib
        h = C.c_void_p()
        _lib.check(self._lib.epde_batch_create(engine._h, *enc.c_args(), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, '_h', None) is not None and self._h:
            self._lib.epde_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:       # noqa: BLE001
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_given(self, mode, wterm, intercept):
        m = np.ascontiguousarray(mode, np.int32)
        w = np.ascontiguousarray(wterm, np.float64)
        b = np.ascontiguousarray(intercept, np.float64)
        assert w.shape == (self.enc.n_eval, MAX_TERMS) and b.shape == (self.enc.n_eval,)
        _lib.check(self._lib.epde_batch_set_given(self._h, _p(m, _lib.c_i32p), _p(w, _lib.c_f64p), _p(b, _lib.c_f64p)))

    def set_norms(self, want_l2=True):
        """which residual norms pass B accumulates (include/epde_b200.h epde_batch_set_norms)."""
        _lib.check(self._lib.epde_batch_set_norms(self._h, int(bool(want_l2))))

    def info(self):
        """{groups, columns, entries, block_pairs, tasks, weighted} of the batch (epde_batch_info)."""
        a = (C.c_int64 * 8)()
        _lib.check(self._lib.epde_batch_info(self._h, a))
        return dict(groups=int(a[0]), columns=int(a[1]), entries=int(a[2]), block_pairs=int(a[3]), tasks=int(a[4]),
                    weighted=bool(a[5]))

    @property
    def gram_size(self):
        r = C.c_int64(0)
        _lib.check(self._lib.epde_batch_gram_size(self._h, C.byref(r)))
        return int(r.value)

    @property
    def residual_size(self):
        r = C.c_int64(0)
        _lib.check(self._lib.epde_batch_residual_size(self._h, C.byref(r)))
        return int(r.value)

    def gram(self, dev_seg=None):
        _lib.check(self._lib.epde_batch_gram(self._h, C.c_void_p(dev_seg) if dev_seg else None))

    def solve(self, dev_seg=None):
        _lib.check(self._lib.epde_batch_solve(self._h, C.c_void_p(dev_seg) if dev_seg else None))

    def solution_size(self, with_sw=False):
        r = C.c_int64(0)
        _lib.check(self._lib.epde_batch_solution_size(self._h, int(bool(with_sw)), C.byref(r)))
        return int(r.value)

    def solve_part(self, dev_seg, part, nparts, dev_solution, with_sw=False):
        """solve 1 / nparts of the evals, zeros elsewhere; solution block -> dev_solution (to be sum-reduced over the ranks)"""
        _lib.check(self._lib.epde_batch_solve_part(self._h, C.c_void_p(dev_seg) if dev_seg else None, int(part), int(nparts),
                                                   int(bool(with_sw)), C.c_void_p(dev_solution)))

    def set_solution(self, dev_solution, with_sw=False):
        _lib.check(self._lib.epde_batch_set_solution(self._h, C.c_void_p(dev_solution), int(bool(with_sw))))

    def residual(self, dev_sums=None):
        _lib.check(self._lib.epde_batch_residual(self._h, C.c_void_p(dev_sums) if dev_sums else None))

    def fetch(self, dev_sums=None, want_sw=True) -> BatchResult:
        o = _alloc_out(self.enc.n_eval, self.engine.n_windows, want_sw)
        _lib.check(self._lib.epde_batch_fetch(self._h, C.c_void_p(dev_sums) if dev_sums else None, *_out_args(o)))
        return _trim(o, self.enc.n_eval)
