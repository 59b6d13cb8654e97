/* epde_b200.h -- C ABI of the B200-native EPDE fitness-evaluation engine.
 *
 * The reference (ITMO-NSS-team/EPDE, pure Python) has no FFI; its boundary for this path is the
 * Python operator / cache API.  Each entry point below names the reference
 * interface it replaces (paths relative to the reference root).  The Python shims in
 * epde_b200/ (ctypes) are the only callers.
 *
 * Conventions: every function returns 0 on success, non-zero on failure (never throws, never
 * aborts); epde_last_error() gives the message of the last failure on the calling thread.
 * Caller owns all host buffers; the library owns device memory unless a parameter is named
 * dev_*.  One CUDA stream per context (epde_ctx_set_stream), every call that returns data to host
 * buffers is synchronous at the call boundary; the staged epde_batch_* calls are asynchronous on
 * the context stream.  One context per process per device (the reference is single threaded with
 * module-global caches: epde/globals.py:26-42).  All arithmetic is IEEE float64.
 *
 * Equation encoding (CSR):  equation e owns terms [term_off[e], term_off[e+1]);  term t owns
 * factors [fac_off[t], fac_off[t+1]);  factor f = (fac_token[f], fac_power[f], fac_kind[f],
 * fac_param[f]):
 *     kind 0: token ** power               epde/evaluators.py:92-135  simple_function_evaluator
 *     kind 1: sin(param * token) ** power  epde/evaluators.py:141-142 trig_eval_fun_np['sin']
 *     kind 2: cos(param * token) ** power  epde/evaluators.py:141     trig_eval_fun_np['cos']
 *     kind 3: token ** (-power)            epde/evaluators.py:147     inverse_eval_fun_np
 * A term is the product of its factors (epde/structure/structure_template.py:49-58,
 * epde/structure/main_structures.py:207-235 with structural=False).
 * An "eval" is one (equation, target term) pair = one LASSOSparsity.apply + L2LRFitness.apply
 * (force_out_of_place=True) of the reference (operators/common/sparsity.py:178-211,
 * operators/common/fitness.py:105-150).
 */
#ifndef EPDE_B200_H
#define EPDE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EPDE_MAX_NDIM 3
#define EPDE_MAX_TERMS 32     /* terms per equation (window systems of up to 32 unknowns: one warp per system) */
#define EPDE_MAX_FACTORS 4    /* factors per term */
#define EPDE_MAX_TOKENS 4096

typedef struct epde_ctx epde_ctx;
typedef struct epde_batch epde_batch;

/* message of the last failing call on this thread ("" if none). */
const char* epde_last_error(void);
/* library / build identification: "epde_b200 <version> sm_100a". */
const char* epde_version(void);

/* ---- token cache ---------------------------------------------------------------------------
 * Replaces epde/cache/cache.py:141-524 (class Cache) for the tensors the evaluator reads, and
 * Cache.set_boundaries / inner_shape (cache.py:250-270): the context is created for one inner
 * grid shape (C order, time axis first), N = prod(inner_shape) points per token vector. */
int epde_ctx_create(int device, int ndim, const int64_t* inner_shape, epde_ctx** out);
int epde_ctx_destroy(epde_ctx* ctx);
/* cudaStream_t to run on (0 = legacy default stream).  torch callers pass
 * torch.cuda.current_stream().cuda_stream so kernels order with NCCL collectives. */
int epde_ctx_set_stream(epde_ctx* ctx, void* cuda_stream);
/* Grid-point sharding: this context holds rows [axis0_offset,
 * axis0_offset + inner_shape[0]) of a global time axis of length axis0_global.  The sliding
 * windows of supplementary.py:411-425 are laid out on the GLOBAL axis.  Default: offset 0,
 * global = inner_shape[0]. */
int epde_ctx_set_shard(epde_ctx* ctx, int64_t axis0_offset, int64_t axis0_global);
/* Cache.add (cache.py:333-395): store / overwrite token `token_id` from a host vector of n = N
 * doubles.  Unlike the reference there is no tensor-count budget: the call fails with an
 * out-of-memory status instead of silently refusing. */
int epde_cache_upload(epde_ctx* ctx, int token_id, const double* host, int64_t n);
/* same, from a device vector: an ASYNCHRONOUS device-to-device copy on the context stream.  Stream contract:
 * the producer of `dev` must have been ordered before the context stream reaches this call (run it on the same
 * stream -- epde_ctx_set_stream -- or synchronise it first); call epde_ctx_synchronize before freeing `dev`. */
int epde_cache_upload_device(epde_ctx* ctx, int token_id, const double* dev, int64_t n);
/* Cache.get (cache.py:397-452): copy the token back to a host vector. */
int epde_cache_download(epde_ctx* ctx, int token_id, double* host, int64_t n);
/* Cache.__contains__ (cache.py:454-470): *present = 1/0. */
int epde_cache_contains(epde_ctx* ctx, int token_id, int* present);
/* Cache.delete_entry (cache.py:472-490). */
int epde_cache_delete(epde_ctx* ctx, int token_id);
/* Cache.consumed_memory (cache.py:305-330): bytes of HBM held by token vectors. */
int epde_cache_bytes(epde_ctx* ctx, int64_t* bytes);
/* sample weights of the window regressions = g_func[g_func_mask] (cache.py:219-244,
 * sparsity.py:202, supplementary.py:455: X^T W X, X^T W y per window).  The LASSO sums, the L1 fitness
 * and the total Gram stay unweighted exactly as in the reference (sparsity.py:54-55,73-76,
 * fitness.py:142).  A vector of ones (the default 0/1 mask restricted to the interior,
 * interface.py:540-545) switches the weighted path off again. */
int epde_ctx_set_sample_weights(epde_ctx* ctx, const double* host, int64_t n);

/* ---- geometry of the sliding windows (supplementary.py:411-425) ----------------------------- */
/* total number of windows W over all axes (rows of calculate_weights' result). */
int epde_ctx_num_windows(epde_ctx* ctx, int* n_windows);
/* total number of segments over all axes (rows of the per-equation segment-Gram table). */
int epde_ctx_num_segments(epde_ctx* ctx, int* n_segments);

/* ---- evaluation -----------------------------------------------------------------------------
 * epde_evaluate_batch = the whole path for a batch of evals, host buffers in and out:
 *   Equation.evaluate (main_structures.py:588-641) -> calculate_weights (supplementary.py:394-471)
 *   -> PhysicsInformedLasso.fit (sparsity.py:42-150) -> LASSOSparsity.apply outputs
 *   (sparsity.py:205-211) -> L2LRFitness.apply (fitness.py:105-174) and L2Fitness.apply
 *   (fitness.py:54-96) values.
 * Outputs, one row per eval, strides in doubles given by the constants above:
 *   coef      [n_eval * EPDE_MAX_TERMS]  weights_internal: feature j of the eval = j-th term of
 *                                         the equation skipping the target (zero padded)
 *   intercept [n_eval]                    PhysicsInformedLasso.intercept_
 *   fitness_l1[n_eval]                    sum|t - Fw - b| / sum|t|   (fitness.py:142); when all
 *                                         coefficients vanish: sum|t - mean(t)| / sum|t|
 *                                         (fitness.py:135-136, 'features is None'; the
 *                                         out-of-place call returns None there: see all_zero)
 *   fitness_l2[n_eval]                    ||Fw + b - t||_2           (fitness.py:78-83, g == 1)
 *   stability [n_eval]                    sum_j CV_j / ndim of the cached window weights
 *                                         (fitness.py:160-169)
 *   sw_weights[n_eval * W * (EPDE_MAX_TERMS+1)] cached_weights_: row w holds n_active[e]+1 values
 *                                         (active features in order, then the intercept), rest 0
 *   n_active  [n_eval]                    number of non-zero coefficients
 *   all_zero  [n_eval]                    1 when every coefficient is zero (fitness.py:123 -> None)
 *   iters     [n_eval]                    (diagnostic) coordinate-descent sweeps executed
 * Any output pointer may be NULL.                                                              */
int epde_evaluate_batch(epde_ctx* ctx,
                        int n_eq, const int32_t* term_off, const int32_t* fac_off,
                        const int32_t* fac_token, const double* fac_power,
                        const int32_t* fac_kind, const double* fac_param,
                        int n_eval, const int32_t* eval_eq, const int32_t* eval_target,
                        double* coef, double* intercept, double* fitness_l1, double* fitness_l2,
                        double* stability, double* sw_weights, int32_t* n_active,
                        int32_t* all_zero, int32_t* iters);

/* Staged form of the same path, for grid-sharded multi-GPU runs: the
 * caller all-reduces (sum, float64) the two device buffers between the stages.
 *   epde_batch_create      upload the encoding
 *   epde_batch_gram        pass A: fused term assembly + per-segment Gram -> dev_seg (device,
 *                          epde_batch_gram_size doubles).  dev_seg == NULL: internal buffer.
 *   epde_batch_solve       window solves + LASSO from (all-reduced) segment Grams
 *   epde_batch_residual    pass B: residual sums -> dev_sums (device, epde_batch_residual_size
 *                          doubles; 3 per eval: sum|r|, sum|t|, sum r^2).  NULL: internal buffer.
 *   epde_batch_fetch       synchronise and copy results to host (same layout as above); reads
 *                          the residual sums from dev_sums (or the internal buffer if NULL).  */
int epde_batch_create(epde_ctx* ctx,
                      int n_eq, const int32_t* term_off, const int32_t* fac_off,
                      const int32_t* fac_token, const double* fac_power,
                      const int32_t* fac_kind, const double* fac_param,
                      int n_eval, const int32_t* eval_eq, const int32_t* eval_target,
                      epde_batch** out);
int epde_batch_destroy(epde_batch* b);
/* In-place fitness of equations whose weights are already known (L2LRFitness.apply /
 * L2Fitness.apply with force_out_of_place=False, fitness.py:70-96,130-174).  Call between
 * epde_batch_create and epde_batch_solve.  Per eval: mode 0 = LASSO (default); 1 = use the given
 * coefficients, still solve the sliding windows on ALL features for the stability objective
 * (fitness.py:155-169 without cached weights); 2 = given coefficients only (stability = NaN).
 * wterm[n_eval * EPDE_MAX_TERMS]: coefficient per TERM index of the equation (target and absent
 * terms 0); intercept[n_eval]. */
int epde_batch_set_given(epde_batch* b, const int32_t* mode, const double* wterm,
                         const double* intercept);
/* Which residual norms pass B accumulates.  want_l2 = 1 (default): sum|r|, sum|t| and sum r^2, i.e. both
 * L2LRFitness (fitness.py:142) and L2Fitness (fitness.py:78-83) values.  want_l2 = 0: the L1 sums only -- what
 * the MOEADD path needs (L2LRFitness is its fitness operator, moeadd/strategy.py:60); fitness_l2 then comes back
 * as NaN.  Call before epde_batch_residual. */
int epde_batch_set_norms(epde_batch* b, int want_l2);
/* Measurement aid: info[0..5] = launch groups, dictionary columns (ones columns included), pair entries,
 * 8x8 column-block pairs pass A computes per 4-point step (= DMMA.8x8x4 per step), pass-A warp tasks, weighted. */
int epde_batch_info(epde_batch* b, int64_t* info);
int epde_batch_gram_size(epde_batch* b, int64_t* n_doubles);
int epde_batch_residual_size(epde_batch* b, int64_t* n_doubles);
int epde_batch_gram(epde_batch* b, double* dev_seg);
int epde_batch_solve(epde_batch* b, const double* dev_seg);
int epde_batch_residual(epde_batch* b, double* dev_sums);
/* Grid-sharded runs need not solve every eval on every rank: epde_batch_solve_part solves the evals at positions
 * part, part + nparts, ... of the batch's solve lists and leaves zeros for the others, then copies the solution block
 * (coefficients, intercepts, stability, ... : epde_batch_solution_size doubles; with_sw = 1 appends the sliding-window
 * weights) to dev_solution.  The caller all-reduces (sum, float64) that buffer over the ranks and hands it back with
 * epde_batch_set_solution before epde_batch_residual. */
int epde_batch_solution_size(epde_batch* b, int with_sw, int64_t* n_doubles);
int epde_batch_solve_part(epde_batch* b, const double* dev_seg, int part, int nparts, int with_sw, double* dev_solution);
int epde_batch_set_solution(epde_batch* b, const double* dev_solution, int with_sw);
int epde_batch_fetch(epde_batch* b, const double* dev_sums,
                     double* coef, double* intercept, double* fitness_l1, double* fitness_l2,
                     double* stability, double* sw_weights, int32_t* n_active,
                     int32_t* all_zero, int32_t* iters);
/* ---- token producers on the device ---------------------------------------
 * epde_fd_gradient = numpy.gradient(data, coords, axis=axis) (edge_order 1) as AdaptiveFiniteDeriv applies it
 * (epde/preprocessing/deriv_calculators.py:46-102; higher orders = repeated calls), on a FULL-grid device
 * tensor of `shape` (C order), result to dev_out (same shape, must differ from dev_in); coords = the host
 * coordinate vector of the axis (shape[axis] values).  Arithmetic is rounded like numpy's (bit-identical).
 * epde_cache_upload_interior = the boundary crop of Cache.set_boundaries / BoundaryExclusion
 * (cache.py:250-270, decorators.py:103-134) + Cache.add: rows [boundary[d], full_shape[d] - boundary[d]) of
 * every axis of a full-grid device tensor become token `token_id` (full - 2 * boundary must equal the
 * context's inner shape). */
/* wait for everything queued on the context stream (the producers above are asynchronous). */
int epde_ctx_synchronize(epde_ctx* ctx);
int epde_fd_gradient(epde_ctx* ctx, const double* dev_in, double* dev_out, int ndim, const int64_t* shape,
                     int axis, const double* coords);
int epde_cache_upload_interior(epde_ctx* ctx, int token_id, const double* dev_full, int ndim,
                               const int64_t* full_shape, const int64_t* boundary);
/* epde_poly_derivative = one derivative of the reference's 'poly' preprocessor (PolynomialDeriv,
 * epde/preprocessing/deriv_calculators.py:111-134 -> process_point_cheb, preprocessing/cheb.py:40-85): the `order`-th
 * derivative along `axis` of the least-squares Chebyshev polynomial through the `window` (odd) values centred on each
 * point; within window / 2 + 1 points of either end the one-sided finite differences of preprocessing/fin_der.py.
 * The fit is linear in the window values: `weights` (host, [shape[axis]][window]) holds, per index of the axis, the
 * derivative-at-the-centre functional as numpy's Chebyshev.fit defines it (computed by epde_b200/producers.py);
 * `coords` = host coordinate vector of the axis.  Full-grid device tensors, dev_out != dev_in, asynchronous. */
int epde_poly_derivative(epde_ctx* ctx, const double* dev_in, double* dev_out, int ndim, const int64_t* shape,
                         int axis, int order, int window, const double* weights, const double* coords);

/* epde_poly_derivative_noalloc: NO COUNTERPART IN THE REFERENCE; added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, checks and result as epde_poly_derivative, but the per-call device scratch (the weight table
 * [shape[axis]][window] and the coordinate vector, which epde_poly_derivative keeps in a cudaMalloc'ed buffer that
 * grows on demand) lives in a fixed file-scope __device__ array of EPDE_NOALLOC_MAX_DOUBLES doubles (512 KiB, held by
 * every GPU that loads the library until the process exits).  The call allocates nothing: no cudaMalloc / cudaFree /
 * allocating library routine.  Limit: shape[axis] * (window + 1) <= EPDE_NOALLOC_MAX_DOUBLES, otherwise the call fails
 * with a message naming the limit (use epde_poly_derivative).  The array is shared by every call on the device and
 * keeps the last call's contents; this call overwrites every value its kernel reads before launching it, so nothing
 * has to be cleared, but two calls whose device work overlaps (different streams / contexts / host threads on one
 * GPU) RACE on it: serialise them.  The first call on a device resolves the array's device address
 * (cudaGetSymbolAddress); like epde_poly_derivative the call copies from pageable host memory and waits for the
 * context stream before it launches. */
#define EPDE_NOALLOC_MAX_DOUBLES 65536
int epde_poly_derivative_noalloc(epde_ctx* ctx, const double* dev_in, double* dev_out, int ndim, const int64_t* shape,
                                 int axis, int order, int window, const double* weights, const double* coords);

/* Measurement aid (no reference counterpart): sustained FP64 rate of `device` in TFLOP/s through plain FMA
 * chains (SASS DFMA) and through the FP64 tensor-core instruction pass A uses (mma.sync m8n8k4 f64, SASS
 * DMMA.8x8x4), each run for about seconds / 2.  The roofline of the streaming passes is FP64 issue, not HBM. */
int epde_measure_fp64_peak(int device, double seconds, double* dfma_tflops, double* dmma_tflops);

/* number of kernel launches issued by the library on this context since creation. */
int epde_ctx_launch_count(epde_ctx* ctx, int64_t* launches);

/* Measurement aid (no reference counterpart): with timing enabled every kernel launch is bracketed
 * by two CUDA events recorded on the context stream; epde_ctx_kernel_time synchronises the stream
 * and returns the summed device time (ms) and launch count of one kernel class since timing was
 * last enabled.  Enabling / disabling resets the sums. */
#define EPDE_KERNEL_GRAM 0       /* pass A: fused term assembly + Gram */
#define EPDE_KERNEL_MARGINAL 1   /* segment / window Gram reductions */
#define EPDE_KERNEL_SOLVE 2      /* window solves + LASSO */
#define EPDE_KERNEL_RESIDUAL 3   /* pass B: residual sums */
#define EPDE_N_KERNEL_CLASSES 4
int epde_ctx_set_timing(epde_ctx* ctx, int enabled);
int epde_ctx_kernel_time(epde_ctx* ctx, int kernel_class, double* ms, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* EPDE_B200_H */
