"""TEST INFRASTRUCTURE ONLY -- CPU (numpy) restatement of EPDE's fitness-evaluation hot path.

Nothing under ``epde_b200/`` may import this module.  It is the *checker* used by
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py``.

Parity status: **pinned against the reference itself run in the build container**
(the reference's own tests hold no numeric vectors for this path):
``oracle/make_golden.py`` imports the unmodified Python reference from ``/root/reference``,
runs its operators on seeded inputs and stores inputs + outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` replays them through this file.

Two restatements live here, each citing the reference lines it follows
(paths relative to ``/root/reference/``):

* ``*_direct``  -- data-space: the same mathematics as the reference evaluated on explicit
  (N, F) matrices with plain slicing instead of ``sliding_window_view``.  Small sizes only.
* ``*_gram``    -- the single-pass formulation the CUDA path implements: every quantity the
  LASSO needs is derived from per-segment Gram matrices of the K = T + 1 columns
  (T term products + ones).  This is what scales to 256^3 / 512^3 where the reference cannot
  allocate its window copies and what serves as the timed CPU port.

Equation encoding used everywhere in this repo (see include/epde_b200.h):
    equation := [term, ...];  term := [factor, ...];
    factor   := (token_id, power, kind, param)
        kind 0  token ** power                 (evaluators.py:92-135 simple_function_evaluator)
        kind 1  sin(param * token) ** power    (evaluators.py:141-142 trig_eval_fun_np['sin'])
        kind 2  cos(param * token) ** power    (evaluators.py:141 trig_eval_fun_np['cos'])
        kind 3  token ** (-power)              (evaluators.py:147 inverse_eval_fun_np)
    ``tokens`` is a list / dict of (N,) float64 vectors (the masked cache entries,
    cache/cache.py:37-63, interface/interface.py:133-155).
"""
from __future__ import annotations

import numpy as np

RIDGE = 1e-6          # supplementary.py:458
MAX_ITER = 20         # sparsity.py:21
TOL = 1e-4            # sparsity.py:21

KIND_POWER, KIND_SIN, KIND_COS, KIND_INVERSE = 0, 1, 2, 3


# ----------------------------------------------------------------------------------------
# window plan  (supplementary.py:411-425)
# ----------------------------------------------------------------------------------------
class AxisPlan:
    """Windows of one axis and the segment decomposition used by the Gram formulation.

    window = n // 2, horizons = window + 1, step = max(1, horizons // 30),
    starts = range(0, horizons, step)  (supplementary.py:412-425).
    ``bounds`` are the sorted distinct window edges plus 0 and n; segment ``s`` covers
    indices [bounds[s], bounds[s+1]).  Window w is the union of segments
    [seg_lo[w], seg_hi[w]).
    """

    def __init__(self, n: int):
        self.n = int(n)
        self.window = self.n // 2
        horizons = self.window + 1
        self.step = max(1, horizons // 30)
        self.starts = np.arange(0, horizons, self.step, dtype=np.int64)
        edges = set([0, self.n])
        for s in self.starts:
            edges.add(int(s))
            edges.add(int(s) + self.window)
        self.bounds = np.array(sorted(edges), dtype=np.int64)
        self.nseg = len(self.bounds) - 1
        pos = {int(b): i for i, b in enumerate(self.bounds)}
        self.seg_lo = np.array([pos[int(s)] for s in self.starts], dtype=np.int64)
        self.seg_hi = np.array([pos[int(s) + self.window] for s in self.starts], dtype=np.int64)
        # segment id of every index along the axis
        self.seg_of = np.repeat(np.arange(self.nseg), np.diff(self.bounds))


class WindowPlan:
    def __init__(self, shape):
        self.shape = tuple(int(s) for s in shape)
        self.axes = [AxisPlan(n) for n in self.shape]
        self.n_windows = int(sum(len(a.starts) for a in self.axes))
        self.n_points = int(np.prod(self.shape))


# ----------------------------------------------------------------------------------------
# term assembly  (factor.py:218-269, evaluators.py:92-151, structure_template.py:49-58,
#                 main_structures.py:207-235 with structural=False)
# ----------------------------------------------------------------------------------------
def eval_factor(tokens, factor):
    tok, power, kind, param = factor
    base = np.asarray(tokens[tok], dtype=np.float64).reshape(-1)
    if kind == KIND_POWER:
        return base if power == 1 else base ** power
    if kind == KIND_SIN:
        return np.sin(param * base) ** power
    if kind == KIND_COS:
        return np.cos(param * base) ** power
    if kind == KIND_INVERSE:
        return np.power(base, -power)
    raise ValueError(f'unknown factor kind {kind}')


def eval_term(tokens, term):
    """reduce(np.multiply, factor values)  (structure_template.py:49-58)."""
    value = eval_factor(tokens, term[0])
    for factor in term[1:]:
        value = value * eval_factor(tokens, factor)
    return value


def term_columns(tokens, equation):
    """(N, T) matrix of all term values of an equation."""
    return np.stack([eval_term(tokens, term) for term in equation], axis=1)


# ----------------------------------------------------------------------------------------
# data-space restatement (small sizes): calculate_weights  (supplementary.py:394-471)
# ----------------------------------------------------------------------------------------
def calculate_weights_direct(X, y, sample_weights, grid_shape):
    X = np.asarray(X, dtype=np.float64)
    n_samples = X.shape[0]
    X_aug = np.hstack([X, np.ones((n_samples, 1))])
    K = X_aug.shape[1]
    plan = WindowPlan(grid_shape)
    Xg = X_aug.reshape(*plan.shape, K)
    yg = np.asarray(y, dtype=np.float64).reshape(plan.shape)
    wg = np.asarray(sample_weights, dtype=np.float64).reshape(plan.shape)
    out = []
    for dim, ax in enumerate(plan.axes):
        for s in ax.starts:
            sl = [slice(None)] * len(plan.shape)
            sl[dim] = slice(int(s), int(s) + ax.window)
            Xw = Xg[tuple(sl)].reshape(-1, K)
            yw = yg[tuple(sl)].reshape(-1)
            ww = wg[tuple(sl)].reshape(-1)
            XTW = Xw.T * ww
            A = XTW @ Xw + RIDGE * np.eye(K)
            b = XTW @ yw
            out.append(np.linalg.solve(A, b))
    return np.vstack(out)


def get_cv(weights):
    """sparsity.py:29-40 (ddof=1, zero-mean -> 0, nan_to_num)."""
    weights = np.asarray(weights, dtype=np.float64)
    if weights.shape[1] == 0:
        return np.zeros(0)
    with np.errstate(divide='ignore', invalid='ignore'):
        std = weights.std(axis=0, ddof=1)
        mu = weights.mean(axis=0)
        cv = (std ** 2) / (mu ** 2)
        cv[mu == 0] = 0.0
    return np.nan_to_num(cv)


def _soft_threshold(x, lam):
    return np.sign(x) * np.maximum(np.abs(x) - lam, 0)      # sparsity.py:26-27


def lasso_fit_direct(X, y, sample_weights, grid_shape, max_iter=MAX_ITER, tol=TOL):
    """PhysicsInformedLasso.fit in data space (sparsity.py:42-150).

    Returns (coef (F,), intercept, cached_weights (W, F'+1)).
    """
    X = np.asarray(X, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n_samples, n_features = X.shape
    weights = calculate_weights_direct(X, y, sample_weights, grid_shape)
    cached = weights
    cv = get_cv(weights[:, :-1])
    coef = weights.mean(axis=0)[:-1].copy()
    intercept = weights.mean(axis=0)[-1]
    norm_sq_features = np.sum(X ** 2, axis=0)
    residual = y - (X @ coef + intercept)
    iteration = 0
    while iteration < max_iter and not np.all(coef == 0):
        max_change = 0
        indices = np.argsort(cv)[::-1]
        for j in indices:
            old_coef = coef[j]
            if old_coef == 0:
                continue
            norm_sq = norm_sq_features[j]
            y_sq_sum = np.sum((y - intercept) ** 2)
            rho = np.dot(X[:, j], residual) + old_coef * norm_sq
            threshold = cv[j] * y_sq_sum
            with np.errstate(divide='ignore', invalid='ignore'):
                new_coef = _soft_threshold(rho, threshold) / norm_sq
            coef[j] = new_coef
            if new_coef == 0:
                mask = coef != 0
                weights = calculate_weights_direct(X[:, mask], y, sample_weights, grid_shape)
                cached = weights
                new_cv = iter(get_cv(weights[:, :-1]))
                cv = np.array([next(new_cv) if m else 0 for m in mask], dtype=np.float64)
                new_c = iter(weights.mean(axis=0)[:-1])
                coef = np.array([next(new_c) if m else 0 for m in mask], dtype=np.float64)
                intercept = weights.mean(axis=0)[-1]
                residual = y - (X @ coef + intercept)
                iteration = 0
                max_change = np.inf
                break
            residual = residual - (new_coef - old_coef) * X[:, j]
            change = abs(new_coef - old_coef)
            if change > max_change:
                max_change = change
        max_abs_coef = np.max(np.abs(coef))
        if max_change <= tol * max_abs_coef:
            xt_residual = X.T @ residual
            y_sq_sum = np.sum((y - intercept) ** 2)
            dual_norm = 0
            for j in range(n_features):
                if cv[j] * y_sq_sum > 0:
                    dual_norm = max(dual_norm, abs(xt_residual[j]) / cv[j] * y_sq_sum)
            const_residual = residual / dual_norm if dual_norm > 1.0 else residual
            primal_obj = 0.5 * np.sum(residual ** 2) + np.sum(cv * y_sq_sum * np.abs(coef))
            dual_obj = 0.5 * y_sq_sum - 0.5 * np.sum((y - intercept - const_residual) ** 2)
            if primal_obj - dual_obj <= tol * (y_sq_sum / n_samples):
                break
        iteration += 1
    return coef, intercept, cached


# ----------------------------------------------------------------------------------------
# Gram formulation (what the CUDA path implements)
# ----------------------------------------------------------------------------------------
def segment_grams(columns, plan: WindowPlan, sample_weights=None):
    """Per-axis per-segment Gram matrices of the K columns (terms + ones).

    columns: (N, T).  Returns list over axes of (nseg_d, K, K) with K = T + 1, ones last.
    Equivalent to X_w^T W X_w of supplementary.py:455-457 summed over the segments of a window.
    """
    N, T = columns.shape
    K = T + 1
    aug = np.empty((N, K))
    aug[:, :T] = columns
    aug[:, T] = 1.0
    grid = aug.reshape(*plan.shape, K)
    wgrid = None if sample_weights is None else np.asarray(sample_weights, np.float64).reshape(plan.shape)
    out = []
    for dim, ax in enumerate(plan.axes):
        Xd = np.moveaxis(grid, dim, 0).reshape(ax.n, -1, K)
        if wgrid is None:
            slab = np.matmul(Xd.transpose(0, 2, 1), Xd)                     # (n_d, K, K)
        else:
            wd = np.moveaxis(wgrid, dim, 0).reshape(ax.n, -1, 1)
            slab = np.matmul(Xd.transpose(0, 2, 1), Xd * wd)
        seg = np.add.reduceat(slab, ax.bounds[:-1], axis=0)
        out.append(seg)
    return out


def window_grams(seg_grams, plan: WindowPlan):
    """(W, K, K): Gram of every sliding window, axes stacked in order (supplementary.py:471)."""
    res = []
    for seg, ax in zip(seg_grams, plan.axes):
        # direct sums over the segments of each window: differences of running sums cancel digits when a column is
        # heavy-tailed along the axis (products of cubed noisy derivatives: 1e-6 relative on a window weight)
        res.append(np.stack([seg[int(lo):int(hi)].sum(axis=0) for lo, hi in zip(ax.seg_lo, ax.seg_hi)], axis=0))
    return np.concatenate(res, axis=0)


def calc_weights_gram(win_grams, feat_idx, y_idx):
    """Ridge solves of supplementary.py:455-469 from window Grams.

    feat_idx: column indices of the active features followed by the ones column.
    """
    idx = np.asarray(feat_idx, dtype=np.int64)
    A = win_grams[:, idx[:, None], idx[None, :]] + RIDGE * np.eye(len(idx))
    b = win_grams[:, idx, y_idx]
    return np.linalg.solve(A, b[..., None])[..., 0]


def window_condition(win_grams, feat_idx):
    """max over windows of cond(D A D), A = ridge system of the columns ``feat_idx``,
    D = diag(A)^-1/2.  Tests use it to tell well-posed evals (result determined by the data) from
    windows that are singular up to the 1e-6 ridge, where the reference's own answer is rounding
    noise of its BLAS summation order (e.g. burgers_b50 eval 5: last x-window, scaled cond 8e15,
    reference and any restatement are both 100 % off the extended-precision solution)."""
    idx = np.asarray(feat_idx, dtype=np.int64)
    A = win_grams[:, idx[:, None], idx[None, :]] + RIDGE * np.eye(len(idx))
    d = 1.0 / np.sqrt(np.abs(np.diagonal(A, axis1=1, axis2=2)))
    return float(np.max(np.linalg.cond(A * d[:, :, None] * d[:, None, :])))


def lasso_fit_gram(win_grams, total_gram, feat_cols, y_col, max_iter=MAX_ITER, tol=TOL):
    """PhysicsInformedLasso.fit (sparsity.py:42-150) driven only by Gram quantities.

    win_grams (W, K, K) (sample-weighted), total_gram (K, K) (unweighted, all N points),
    feat_cols: the F feature column indices (terms other than the target, in term order),
    y_col: target column.  The ones column is K-1.  Returns (coef, intercept, cached_weights).
    """
    K = total_gram.shape[0]
    ones = K - 1
    fc = np.asarray(feat_cols, dtype=np.int64)
    F = len(fc)
    G = total_gram[np.ix_(fc, fc)]             # X^T X
    Xy = total_gram[fc, y_col]                 # X^T y
    Xs = total_gram[fc, ones]                  # X^T 1
    yy = total_gram[y_col, y_col]
    sy = total_gram[y_col, ones]
    N = total_gram[ones, ones]

    def refit(mask):
        act = list(fc[mask]) + [ones]
        w = calc_weights_gram(win_grams, act, y_col)
        cv_act = iter(get_cv(w[:, :-1]))
        mu = w.mean(axis=0)
        c_act = iter(mu[:-1])
        cv_full = np.array([next(cv_act) if m else 0 for m in mask], dtype=np.float64)
        c_full = np.array([next(c_act) if m else 0 for m in mask], dtype=np.float64)
        return w, cv_full, c_full, mu[-1]

    cached, cv, coef, intercept = refit(np.ones(F, dtype=bool))
    norm_sq_features = np.diag(G).copy()

    def xt_res(j=None):
        # X_j^T (y - X c - b)
        if j is None:
            return Xy - G @ coef - intercept * Xs
        return Xy[j] - G[j] @ coef - intercept * Xs[j]

    def ysq():
        return yy - 2.0 * intercept * sy + N * intercept * intercept

    iteration = 0
    while iteration < max_iter and not np.all(coef == 0):
        max_change = 0
        indices = np.argsort(cv, kind='stable')[::-1]
        for j in indices:
            old_coef = coef[j]
            if old_coef == 0:
                continue
            norm_sq = norm_sq_features[j]
            y_sq_sum = ysq()
            rho = xt_res(j) + old_coef * norm_sq
            threshold = cv[j] * y_sq_sum
            with np.errstate(divide='ignore', invalid='ignore'):
                new_coef = _soft_threshold(rho, threshold) / norm_sq
            coef[j] = new_coef
            if new_coef == 0:
                cached, cv, coef, intercept = refit(coef != 0)
                iteration = 0
                max_change = np.inf
                break
            change = abs(new_coef - old_coef)
            if change > max_change:
                max_change = change
        max_abs_coef = np.max(np.abs(coef))
        if max_change <= tol * max_abs_coef:
            xr = xt_res()
            y_sq_sum = ysq()
            dual_norm = 0
            for j in range(F):
                if cv[j] * y_sq_sum > 0:
                    dual_norm = max(dual_norm, abs(xr[j]) / cv[j] * y_sq_sum)
            # r = y - Xc - b ;  all sums below follow from the Gram
            cGc = coef @ G @ coef
            cXy = coef @ Xy
            cXs = coef @ Xs
            rr = yy - 2 * cXy - 2 * intercept * sy + cGc + 2 * intercept * cXs + N * intercept ** 2
            sum_r = sy - cXs - N * intercept
            y_r = yy - cXy - intercept * sy
            yb_r = y_r - intercept * sum_r                       # (y - b)^T r
            alpha = 1.0 / dual_norm if dual_norm > 1.0 else 1.0
            primal_obj = 0.5 * rr + np.sum(cv * y_sq_sum * np.abs(coef))
            dual_obj = 0.5 * y_sq_sum - 0.5 * (y_sq_sum - 2 * alpha * yb_r + alpha * alpha * rr)
            if primal_obj - dual_obj <= tol * (y_sq_sum / N):
                break
        iteration += 1
    return coef, intercept, cached


def stability_from_weights(weights, ndim):
    """fitness.py:160-169: sum(cv[:-1]) / ndim with ddof=1 and cv:=0 where mean==0 (no nan_to_num)."""
    w = np.asarray(weights, dtype=np.float64)
    with np.errstate(divide='ignore', invalid='ignore'):
        std = w.std(axis=0, ddof=1)
        mu = w.mean(axis=0)
        cv = (std ** 2) / (mu ** 2)
        cv[mu == 0] = 0.0
    return float(sum(cv[:-1]) / ndim)


def l1_fitness(target, features, coef_nonzero, intercept):
    """fitness.py:135-142: sum|t - (F w + b)| / sum|t|; features None -> t - mean(t)."""
    if features is None or features.shape[1] == 0:
        discr = target - target.mean()
    else:
        discr = target - (features @ np.asarray(coef_nonzero) + intercept)
    return float(np.sum(np.abs(discr)) / np.sum(np.abs(target)))


def l2_fitness(target, features, coef_nonzero, intercept, g_flat=None, penalty_coeff=0.2,
               weights_final=None):
    """fitness.py:70-91 (L2Fitness): ||(F w + b - t) * g||_2, / penalty if sum(weights_final)==0."""
    feats = 0 if features is None or features.shape[1] == 0 else features @ np.asarray(coef_nonzero)
    discr = feats + intercept - target
    if g_flat is not None:
        discr = discr * g_flat
    val = float(np.linalg.norm(discr, ord=2))
    if weights_final is not None and np.sum(weights_final) == 0:
        val /= penalty_coeff
    return val


# ----------------------------------------------------------------------------------------
# whole path for one equation
# ----------------------------------------------------------------------------------------
class EvalResult:
    """What the reference leaves on the Equation after LASSOSparsity + L2LRFitness
    (sparsity.py:205-211, fitness.py:142-174)."""
    __slots__ = ('target', 'coef', 'intercept', 'sw_weights', 'fitness', 'fitness_l2',
                 'stability', 'all_zero', 'cond')

    def as_dict(self):
        return {k: getattr(self, k) for k in self.__slots__}


def evaluate_equation(tokens, grid_shape, equation, targets=None, sample_weights=None,
                      method='gram', columns=None):
    """Every admissible target of one equation through LASSO + L1 fitness (+ stability).

    Mirrors the body of EqRightPartSelector's loop (right_part_selection.py:60-74):
    L2LRFitness.apply(force_out_of_place=True) per target, plus the quantities the later
    in-place call adds (fitness.py:152-174).  ``method``: 'gram' or 'direct'.
    """
    plan = WindowPlan(grid_shape)
    cols = term_columns(tokens, equation) if columns is None else columns
    N, T = cols.shape
    if targets is None:
        targets = list(range(T))
    g = np.ones(N) if sample_weights is None else np.asarray(sample_weights, np.float64)
    if method == 'gram':
        seg_w = segment_grams(cols, plan, None if sample_weights is None else g)
        win = window_grams(seg_w, plan)
        if sample_weights is None:
            total = seg_w[0].sum(axis=0)
        else:
            total = segment_grams(cols, plan, None)[0].sum(axis=0)
    results = []
    for tgt in targets:
        feat = [i for i in range(T) if i != tgt]
        if method == 'gram':
            coef, b, cached = lasso_fit_gram(win, total, feat, tgt)
        else:
            coef, b, cached = lasso_fit_direct(cols[:, feat], cols[:, tgt], g, plan.shape)
        r = EvalResult()
        r.cond = window_condition(win, feat + [T], ) if method == 'gram' else float('nan')
        r.target = tgt
        r.coef = coef
        r.intercept = float(b)
        r.sw_weights = cached
        r.all_zero = bool(np.all(coef == 0))
        nz = coef != 0
        fsel = cols[:, np.asarray(feat, dtype=np.int64)[nz]] if len(feat) else None
        # weights_final strips zeros incl. a zero intercept (sparsity.py:207-209); the value used
        # as intercept by fitness.py:138-139 is weights_final[-1].
        r.fitness = l1_fitness(cols[:, tgt], fsel, coef[nz], b)
        r.fitness_l2 = l2_fitness(cols[:, tgt], fsel, coef[nz], b, g)
        r.stability = stability_from_weights(cached, len(plan.shape))
        results.append(r)
    return results


def select_right_part(results):
    """argmin over targets with fitness not None (all-zero -> None), strict '<'
    (right_part_selection.py:66-71).  Returns index into ``results`` or None."""
    best, best_fit = None, np.inf
    for i, r in enumerate(results):
        if r.all_zero:
            continue
        if r.fitness < best_fit:
            best, best_fit = i, r.fitness
    return best


# ----------------------------------------------------------------------------------------
# streamed variant for large grids (256^3 / 512^3): same Gram formulation, axis-0 chunks so that
# host memory stays bounded.  Used as the timed CPU port where the reference cannot allocate its
# window copies and by the large-size spot checks.
# ----------------------------------------------------------------------------------------
def _chunk_rows(plan, T, budget_bytes=1 << 30):
    row = int(np.prod(plan.shape[1:])) if len(plan.shape) > 1 else 1
    return max(1, int(budget_bytes // (8 * row * (T + 2) * 2)))


def segment_grams_streamed(tokens, equation, plan: WindowPlan, rows_per_chunk=None):
    """segment_grams() without materialising the (N, K) matrix: loops over chunks of axis 0."""
    T = len(equation)
    K = T + 1
    shape = plan.shape
    nd = len(shape)
    n0 = shape[0]
    row = int(np.prod(shape[1:])) if nd > 1 else 1
    if rows_per_chunk is None:
        rows_per_chunk = _chunk_rows(plan, T)
    slab0 = np.zeros((n0, K, K))
    others = [np.zeros((shape[d], K, K)) for d in range(1, nd)]
    for lo in range(0, n0, rows_per_chunk):
        hi = min(n0, lo + rows_per_chunk)
        sub = [np.asarray(t).reshape(-1)[lo * row:hi * row] for t in tokens] if not isinstance(tokens, dict) \
            else {k: np.asarray(v).reshape(-1)[lo * row:hi * row] for k, v in tokens.items()}
        aug = np.empty(((hi - lo) * row, K))
        for t, term in enumerate(equation):
            aug[:, t] = eval_term(sub, term)
        aug[:, T] = 1.0
        grid = aug.reshape((hi - lo,) + tuple(shape[1:]) + (K,))
        X0 = grid.reshape(hi - lo, -1, K)
        slab0[lo:hi] = np.matmul(X0.transpose(0, 2, 1), X0)
        for d in range(1, nd):
            Xd = np.moveaxis(grid, d, 0).reshape(shape[d], -1, K)
            others[d - 1] += np.matmul(Xd.transpose(0, 2, 1), Xd)
    slabs = [slab0] + others
    return [np.add.reduceat(s, ax.bounds[:-1], axis=0) for s, ax in zip(slabs, plan.axes)]


def residual_sums_streamed(tokens, equation, plan, jobs, rows_per_chunk=None):
    """jobs: [(target, coef_per_term (T,), b_l1, b_l2)] -> [(sum|r|, sum|t|, sum r2^2)]."""
    T = len(equation)
    shape = plan.shape
    row = int(np.prod(shape[1:])) if len(shape) > 1 else 1
    if rows_per_chunk is None:
        rows_per_chunk = _chunk_rows(plan, T)
    acc = np.zeros((len(jobs), 3))
    for lo in range(0, shape[0], rows_per_chunk):
        hi = min(shape[0], lo + rows_per_chunk)
        sub = [np.asarray(t).reshape(-1)[lo * row:hi * row] for t in tokens]
        cols = np.stack([eval_term(sub, term) for term in equation], axis=1)
        for i, (tgt, w, b1, b2) in enumerate(jobs):
            pred = cols @ w
            tcol = cols[:, tgt]
            acc[i, 0] += np.abs(tcol - (pred + b1)).sum()
            acc[i, 1] += np.abs(tcol).sum()
            acc[i, 2] += ((pred + b2 - tcol) ** 2).sum()
    return acc


def evaluate_equation_streamed(tokens, grid_shape, equation, targets=None):
    """evaluate_equation(method='gram') for grids too large for the in-memory version."""
    plan = WindowPlan(grid_shape)
    T = len(equation)
    if targets is None:
        targets = list(range(T))
    seg = segment_grams_streamed(tokens, equation, plan)
    win = window_grams(seg, plan)
    total = seg[0].sum(axis=0)
    results, jobs = [], []
    for tgt in targets:
        feat = [i for i in range(T) if i != tgt]
        coef, b, cached = lasso_fit_gram(win, total, feat, tgt)
        r = EvalResult()
        r.cond = window_condition(win, feat + [T])
        r.target, r.coef, r.intercept, r.sw_weights = tgt, coef, float(b), cached
        r.all_zero = bool(np.all(coef == 0))
        r.stability = stability_from_weights(cached, len(plan.shape))
        w = np.zeros(T)
        w[feat] = coef
        b1 = total[tgt, T] / total[T, T] if r.all_zero else float(b)      # fitness.py:135-136
        jobs.append((tgt, w, b1, float(b)))
        results.append(r)
    sums = residual_sums_streamed(tokens, equation, plan, jobs)
    for r, s in zip(results, sums):
        r.fitness = float(s[0] / s[1])
        r.fitness_l2 = float(np.sqrt(s[2]))
    return results


# ----------------------------------------------------------------------------------------
# token producer: the 'poly' preprocessor (preprocessing/deriv_calculators.py:111-134, cheb.py:14-85, fin_der.py)
# restated point by point for SMALL grids (pure Python loop over points, as the reference does in its process pool)
# ----------------------------------------------------------------------------------------
def poly_derivative_points(data, coords, axis, order, n_der, window=9, poly_order=None):
    """order-th derivative along `axis` of `data` (full grid) exactly as process_point_cheb computes it per point:
    Chebyshev.fit over the window at interior indices (cheb.py:14-37), one-sided finite differences within
    window // 2 + 1 points of the ends (fin_der.py:12-36).  coords: 1-D coordinate vector of the axis."""
    from numpy.polynomial.chebyshev import Chebyshev
    data = np.asarray(data, dtype=np.float64)
    x = np.asarray(coords, dtype=np.float64).reshape(-1)
    n = data.shape[axis]
    bound = window // 2 + 1
    deg = (n_der + 1) if poly_order is None else poly_order
    if window == deg:
        deg -= 1
    moved = np.moveaxis(data, axis, 0).reshape(n, -1)
    out = np.empty_like(moved)
    h = (window - 1) // 2
    for i in range(n):
        if bound <= i <= n - bound:
            xr = x[i - h:i + h + 1]
            for c in range(moved.shape[1]):
                out[i, c] = Chebyshev.fit(xr, moved[i - h:i + h + 1, c], deg).deriv(m=order)(xr[int(xr.size / 2.)])
        else:
            I = i + np.arange(6) if i < bound else i - np.arange(6)
            F, xx = moved[I], x[I]
            dx = xx[1] - xx[0]
            if order == 1:
                out[i] = (F[1] - F[0]) / dx
            elif order == 2:
                out[i] = (2 * F[0] - 5 * F[1] + 4 * F[2] - F[3]) / dx ** 2
            else:
                out[i] = (-2.5 * F[0] + 9 * F[1] - 12 * F[2] + 7 * F[3] - 1.5 * F[4]) / dx ** 3
    shape = (n,) + tuple(s for d, s in enumerate(data.shape) if d != axis)
    return np.moveaxis(out.reshape(shape), 0, axis)
