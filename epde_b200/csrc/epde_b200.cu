// epde_b200.cu -- B200 (sm_100a) kernels + C ABI of the EPDE fitness-evaluation engine.
//
// Path (reference paths relative to the EPDE tree):
//   token cache            epde/cache/cache.py:141-524              -> HBM-resident token table
//   term assembly          structure/factor.py:218-269, evaluators.py:92-151,
//                          structure/main_structures.py:207-235,588-641
//                                                                    -> build_columns (both passes)
//   calculate_weights      epde/supplementary.py:394-471             -> gram_dmma_kernel (pass A), marginal_*,
//                                                                       window_gram_kernel, lasso_solve_kernel
//   PhysicsInformedLasso   operators/common/sparsity.py:20-150       -> lasso_solve_kernel
//   L2LRFitness/L2Fitness  operators/common/fitness.py:27-183        -> residual_cols_kernel (pass B)
//
// Design (round 2): the equations of a launch group are de-duplicated into a dictionary of DISTINCT term
// columns (a population of 128 equations with 630 term instances holds 68 distinct terms) plus a ones column.
// Both streaming passes first build, per tile of points, the value of every distinct column ONCE in shared
// memory (phase 1), then
//   pass A  accumulates the Gram of the column dictionary with the FP64 tensor-core instruction
//           (mma.sync m8n8k4 f64 = SASS DMMA.8x8x4): 8x8 column-block pairs, fragments read straight from the
//           column tile, accumulators in registers, one warp per (<= 2 x 3 block pairs) task.  The DMMA pipe
//           is the FP64 pipe (measured: 37.1 TFLOP/s DMMA vs 34 DFMA, not additive), but one DMMA does the
//           work of eight warp-wide DFMAs with two operand registers, so issue slots, register bandwidth and
//           shared-memory traffic stop being the limit.  Only the pair entries some equation needs are
//           written to the per-cell table.
//   pass B  evaluates each (equation, target) residual as a sparse dot over its target + surviving columns.
// Every equation then addresses the shared pair table through its column list, so a Gram entry is computed
// once per launch group however many equations use it, and results do not depend on batch composition
// beyond the grouping itself.
//
// Data layout in HBM
//   token t            double[N]   N = prod(inner_shape), C order, last axis contiguous
//   cell table  C      per launch group: [cellA][cellB][piece][n_ent] -- n_ent = needed (column, column) pair sums
//   segment table S    per launch group: [segment row][n_ent] (the (ones, ones) entry = point count / weight sum);
//                      rows of all axes stacked.  This is the buffer that is all-reduced under grid sharding.
//   window table  Wg   per launch group: [W + 1][n_ent]: row w = sum of the segment rows of sliding window w
//                      (direct summation, no prefix differences); row W = all points (unweighted).
//
// No CPU fallback exists in this file: every entry point either runs the CUDA path or fails.
#include "../../include/epde_b200.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#define TMAX EPDE_MAX_TERMS
#define FMAXF EPDE_MAX_FACTORS
#define KS (TMAX + 1)          // stride of a sliding-window weight row
#define RIDGE 1e-6             // supplementary.py:458
#define LASSO_MAX_ITER 20      // sparsity.py:21
#define LASSO_TOL 1e-4         // sparsity.py:21

#define MAXU 32                // distinct tokens per launch group (the shared-memory budget may bind first)
#define NBLK_MAX 10            // 8-column blocks of the column dictionary of a launch group
#define CMAX (NBLK_MAX * 8)    // columns per launch group, ones column included
#define MAXSF 6                // token values multiplied by a "fast" column (integer powers expanded)
#define TILE_ROWS 64           // rows (leading-axes index pairs) of a pass-A unit (its width is chosen per grid)
#define PT_A 128               // points per pass-A chunk = columns of the column tile
#define PP_A (PT_A + 4)        // its pitch: == 4 (mod 16) doubles -> conflict-free DMMA fragment loads
#define TASK_A 2               // a pass-A task covers TASK_A x TASK_B block pairs
#define TASK_B 3
#define GRAM_MAX_WARPS 12     // consumer (tensor-core) warps of pass A
#ifndef GRAM_PROD_WARPS
#define GRAM_PROD_WARPS 10     // producer (term assembly) warps of pass A (fewer when the consumers need the slots)
#endif
#define GRAM_MAX_THREADS 704
#define PT_B 256               // points per pass-B chunk
#define PP_B (PT_B + 4)
#define RES_WARPS 16

static thread_local std::string g_err;

static int fail(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t _e = (call);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// ------------------------------------------------------------------------------------------
// device-side descriptors
// ------------------------------------------------------------------------------------------
struct FactorDesc {          // 32 bytes; factors of the generic (non-fast) columns of a launch group
    int slot;                // token slot inside the launch group's tile
    int ipow;                // the power if it is an integer in 1..64 (multiplications, no pow()), else 0
    int kind;                // 0 pow, 1 sin, 2 cos, 3 inverse
    int pad;
    double power;
    double param;            // freq for trig kinds
};

struct ColDesc {             // 32 bytes; one distinct term of a launch group = one column of the dictionary
    int off[MAXSF];          // fast column: token slots of the expanded factors (u^2 * v -> u, u, v), -1 = none; the
                             // kernels scale them to byte offsets of the slot inside a token tile when staging
    short nf;                // number of expanded factors (fast) / factors (generic)
    short kind;              // 0 fast, 1 generic, 2 ones column, 3 padding (all zero), 4 product of <= 4 integer powers
    int gfac_off;            // generic: first factor in the group's factor array; kind 4: the powers, 8 bits each
};
static_assert(sizeof(ColDesc) == 32, "ColDesc is copied as int4 pairs");

struct EqDesc {
    int T;
    int group;               // launch group
    int ones_col;            // column index of the ones column in the group's dictionary
    int n_ent;               // pair entries of the group = row length of its tables
    long long w_off;         // doubles, into the window table (start of the group's table)
    long long pe_off;        // ints, into the pair -> entry lookup tables (start of the group's CP x CP table)
    int cp;                  // padded column count of the group (row length of the lookup table)
    int pad;
    short col[TMAX + 1];     // term index -> column of the dictionary
};

struct UnitDesc {            // one CTA work unit of pass A: a (sub-)cell of the leading axes x a column chunk
    int a_lo, a_hi;          // index range on leading slot A
    int b_lo, b_hi;          // index range on leading slot B
    int col_lo, ncols;       // columns (last axis) handled, ncols <= TILE_COLS
    int cell;                // (cellA * nB + cellB)
    int piece_lo, npieces;   // pieces (sub-segments of the last axis) covered by the chunk
    int pad0, pad1, pad2;
};

struct MargDesc {            // one output row of the segment table
    int row;                 // row in S
    int a_lo, a_hi, b_lo, b_hi, p_lo, p_hi;   // cell ranges to sum
    int kind;                // which range is the segment's: 0 = slot A, 1 = slot B, 2 = pieces
    double count;            // number of grid points in the segment slab (local)
};

struct Geom {                // kernel-visible geometry (by value)
    long long strideA, strideB;    // element strides of the leading slots
    int nA, nB, npieces;           // cell-grid extents
    int n_units;
    int n_cells;                   // nA * nB * npieces
    int n_rows;                    // segment rows over all axes
    int n_windows;
    int ndim;
    int max_rows;                  // largest row count of a unit (<= TILE_ROWS)
    int aligned16;                 // 1: every tile row can be fetched with 16-byte aligned bulk copies
    long long N;
    const UnitDesc* units;
    const int* piece_start;        // [npieces + 1] column bounds
    const MargDesc* marg;          // [n_rows]
    const int* win_lo_row;         // [W + 1] first segment row of window w (entry W: all of axis 0)
    const int* win_hi_row;         // [W + 1] one past the last segment row
};

// ------------------------------------------------------------------------------------------
// shared-memory / async-copy helpers (sm_100a: bulk async copies completing on an mbarrier)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    const unsigned a = smem_addr(bar);
    unsigned done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    }
}
// global -> shared bulk copy (TMA engine, SASS UBLKCP); bytes multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ double lds_f64(unsigned a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}

// FP64 tensor-core MMA: D(8x8) += A(8x4, row) * B(4x8, col).  Lane l holds A[l/4][l%4], B[l%4][l/4] and
// D[l/4][2*(l%4) + {0,1}].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// x / d for 0 <= x < 2^24, d >= 1, through a float reciprocal and one correction step (the producers of pass A locate
// 4 points per lane and chunk: two runtime integer divisions each were 14 % of the kernel's instructions)
__device__ __forceinline__ int small_div(int x, int d, float inv_d) {
    int q = __float2int_rz((float)x * inv_d);
    const int rem = x - q * d;
    q += (rem >= d) ? 1 : 0;
    q -= (rem < 0) ? 1 : 0;
    return q;
}

// ------------------------------------------------------------------------------------------
// factor / term evaluation (factor.py:218-269, evaluators.py:92-151)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double apply_factor(const FactorDesc& f, double x) {
    if (f.kind != 0) {
        if (f.kind == 1) x = sin(f.param * x);
        else if (f.kind == 2) x = cos(f.param * x);
        else return pow(x, -f.power);
    }
    if (f.ipow == 1) return x;
    if (f.ipow == 2) return x * x;
    if (f.ipow == 3) return x * x * x;
    if (f.ipow > 3) {                        // integer power by squaring (uniform per warp: a column is one warp's job)
        double r = (f.ipow & 1) ? x : 1.0, sq = x;
        for (int e = f.ipow >> 1; e; e >>= 1) {
            sq *= sq;
            if (e & 1) r *= sq;
        }
        return r;
    }
    return pow(x, f.power);
}

// Phase 1 of both passes: the value of every column of the launch group's dictionary at the points of the
// current chunk -> column tile ct[c][p] (shared memory, row pitch ppb bytes).  Warps walk over columns (a column
// is uniform per warp: no divergence), lanes over points: thread handles points p = lane + 32 k, k < NPL, whose
// token values sit at tile + (slot offset) + toffb[k]; invalid (padding) points get 0 in EVERY column, the ones
// column included, so they add nothing to any sum.  All operands are addressed in the shared window explicitly
// (32-bit addresses, ld.shared / st.shared): no generic-pointer arithmetic in the inner loops.
// s_cols holds the ColDesc of the group with `off` already scaled to byte offsets of the slot inside a tile.
__device__ __forceinline__ int4 lds_v4(unsigned a) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
// shared load the compiler may schedule freely among the arithmetic of its consumers: not volatile, but tied to an
// `epoch` value that is produced after the barrier which made the data visible (so it cannot be hoisted above that
// barrier) and that differs between jobs (so loads of different jobs are never merged)
__device__ __forceinline__ double lds_f64_e(unsigned a, int epoch) {
    double v;
    asm("ld.shared.f64 %0, [%1];  // epoch %2" : "=d"(v) : "r"(a), "r"(epoch));
    return v;
}
__device__ __forceinline__ void sts_f64(unsigned a, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory");
}

__device__ __forceinline__ void load_cols_to_smem(ColDesc* s_cols, const ColDesc* cols, int nbuild, int tstride_b,
                                                  int tid, int nthreads) {
    for (int q = tid; q < nbuild; q += nthreads) {
        ColDesc cd = cols[q];
#pragma unroll
        for (int f = 0; f < MAXSF; ++f) cd.off[f] = (cd.off[f] >= 0) ? cd.off[f] * tstride_b : 0;
        s_cols[q] = cd;
    }
}

template <int NPL, bool LISTED = false>
__device__ __forceinline__ void build_columns(unsigned ct, int ppb, unsigned cols, int nbuild, const FactorDesc* gfac,
                                              unsigned tile, int tstride_b, const int (&toffb)[NPL],
                                              const bool (&valid)[NPL], int warp, int nwarps, int lane, int epoch,
                                              unsigned list = 0) {
    // LISTED: `list` = shared-memory address of the nbuild column numbers to build (pass B builds only the columns
    // some eval reads); otherwise all columns 0 .. nbuild - 1
    for (int ci = warp; ci < nbuild; ci += nwarps) {
        int c = ci;
        if (LISTED) asm volatile("ld.shared.s32 %0, [%1];" : "=r"(c) : "r"(list + ci * 4));
        const int4 q0 = lds_v4(cols + c * 32), q1 = lds_v4(cols + c * 32 + 16);
        const int nf = q1.z & 0xffff, kind = q1.z >> 16;
        double v[NPL];
        if (kind == 0) {
#pragma unroll
            for (int k = 0; k < NPL; ++k) v[k] = lds_f64_e(tile + q0.x + toffb[k], epoch);
            if (nf > 1) {
#pragma unroll
                for (int k = 0; k < NPL; ++k) v[k] *= lds_f64_e(tile + q0.y + toffb[k], epoch);
                if (nf > 2) {
#pragma unroll
                    for (int k = 0; k < NPL; ++k) v[k] *= lds_f64_e(tile + q0.z + toffb[k], epoch);
                    if (nf > 3) {
#pragma unroll
                        for (int k = 0; k < NPL; ++k) v[k] *= lds_f64_e(tile + q0.w + toffb[k], epoch);
                        if (nf > 4) {
#pragma unroll
                            for (int k = 0; k < NPL; ++k) v[k] *= lds_f64_e(tile + q1.x + toffb[k], epoch);
                            if (nf > 5) {
#pragma unroll
                                for (int k = 0; k < NPL; ++k) v[k] *= lds_f64_e(tile + q1.y + toffb[k], epoch);
                            }
                        }
                    }
                }
            }
        } else if (kind == 1) {
#pragma unroll
            for (int k = 0; k < NPL; ++k) v[k] = 1.0;
            for (int f = 0; f < nf; ++f) {
                const FactorDesc fd = gfac[q1.w + f];
#pragma unroll
                for (int k = 0; k < NPL; ++k) v[k] *= apply_factor(fd, lds_f64_e(tile + fd.slot * tstride_b + toffb[k], epoch));
            }
        } else if (kind == 4) {
            // product of up to 4 token values raised to integer powers that do not fit the expanded form (u^5 * v^4):
            // slots in off[0..3], powers in gfac_off; powers by squaring, everything warp-uniform, no global loads
#pragma unroll
            for (int k = 0; k < NPL; ++k) v[k] = 1.0;
            for (int f = 0; f < nf; ++f) {
                const int o = (f == 0) ? q0.x : (f == 1) ? q0.y : (f == 2) ? q0.z : q0.w;
                const int pw = (q1.w >> (8 * f)) & 0xff;
                double x[NPL];
#pragma unroll
                for (int k = 0; k < NPL; ++k) x[k] = lds_f64_e(tile + o + toffb[k], epoch);
                if (pw & 1) {
#pragma unroll
                    for (int k = 0; k < NPL; ++k) v[k] *= x[k];
                }
                for (int e = pw >> 1; e; e >>= 1) {
#pragma unroll
                    for (int k = 0; k < NPL; ++k) x[k] *= x[k];
                    if (e & 1) {
#pragma unroll
                        for (int k = 0; k < NPL; ++k) v[k] *= x[k];
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < NPL; ++k) v[k] = (kind == 2) ? 1.0 : 0.0;
        }
        const unsigned dst = ct + c * ppb + lane * 8;
#pragma unroll
        for (int k = 0; k < NPL; ++k) sts_f64(dst + k * 256, valid[k] ? v[k] : 0.0);
    }
}

// ------------------------------------------------------------------------------------------
// pass A: Gram of the column dictionary per (cell, piece), FP64 tensor cores.
//
// Persistent CTAs (two per SM) walk the units (a (sub-)cell of the leading axes x a chunk of <= tile_cols
// last-axis columns).  The token tile of a unit is fetched by the TMA engine: one tensor copy per token
// (cp.async.bulk.tensor, box = unit rows x unit columns of the token's N-d tensor map, SASS UTMALDG) completing
// on an mbarrier; the second CTA of the SM computes while this one waits.  Every last-axis piece of the unit is a
// RUN of rows x piece-width points with its own Gram; a run is processed in chunks of PT_A points:
//   phase 1  all warps build the column tile ct[column][point] (build_columns),
//   phase 2  warp (task, kpart) accumulates its <= TASK_A x TASK_B block pairs with DMMA.8x8x4 over the 4-point
//            k-steps s == kpart (mod ksplit) of the chunk; accumulators stay in registers across the chunks of a run.
// At the end of a run the D fragments are scattered to the cell record through the task's flush map (pair entry
// index per (pair slot, lane, half), -1 = not needed); with ksplit > 1 the partial fragments of the kparts are first
// summed in shared memory in kpart order.  No atomics: every table entry has one writer and a fixed summation order.
// ------------------------------------------------------------------------------------------
struct TaskDesc {            // 8 bytes
    unsigned char ablk[TASK_A], bblk[TASK_B];
    unsigned char mask;      // bit (i * TASK_B + j): block pair (ablk[i], bblk[j]) belongs to this task
    unsigned char pad0, pad1;
};
#define TASK_PAIRS (TASK_A * TASK_B)

struct GramParams {
    Geom g;
    const ColDesc* cols;     // [nbuild]: dictionary columns, zero padding up to cp, then the weight pseudo column
    const FactorDesc* gfac;
    int cp;                  // 8 * number of column blocks
    int nbuild;              // columns phase 1 builds = rows of the column tile (cp, + 1 when weighted)
    int wcol;                // weighted mode: row of the column tile holding the sample weight of each point
    int U;                   // token slots staged per tile (weight token included)
    const double* tok[MAXU + 1];
    const TaskDesc* tasks;
    int ntasks, ksplit;
    const int* flushmap;     // [ntasks][TASK_PAIRS][32][2]
    double* cells;           // [n_cells][n_ent]
    int n_ent;
    int use_tma;             // 1: token tiles come through the tensor maps below; 0: plain loads (odd last axis)
    int nstages;             // token-tile buffers (2: the next unit's tile lands while this one is consumed)
    int box_a, box_b, box_c; // tile extents = TMA box (slot A rows, slot B rows, last-axis columns)
    alignas(64) CUtensorMap tmap[MAXU + 1];
};

__device__ __forceinline__ void tma_load_tile(double* dst, const CUtensorMap* tm, int ndim, int c0, int c1, int c2, uint64_t* bar) {
    const unsigned d = smem_addr(dst), b = smem_addr(bar);
    if (ndim == 3)
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(d), "l"(tm), "r"(b), "r"(c0), "r"(c1), "r"(c2) : "memory");
    else if (ndim == 2)
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(d), "l"(tm), "r"(b), "r"(c0), "r"(c1) : "memory");
    else
        asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3}], [%2];"
                     ::"r"(d), "l"(tm), "r"(b), "r"(c0) : "memory");
}

// the k-step loop of one consumer warp, specialised on the task's pair mask: a predicated-off DMMA occupies the FP64
// tensor pipe exactly like an executed one (measured: scratch/dmma_feed.cu), so inactive pair slots must not exist in
// the instruction stream at all -- one instantiation per mask, selected once per chunk.
template <int MASK, bool WEIGHTED>
__device__ __forceinline__ void gram_ksteps(double (&acc)[TASK_PAIRS][2], unsigned ct, const unsigned (&a_off)[TASK_A],
                                            const unsigned (&b_off)[TASK_B], unsigned w_off, int kpart, int nks, int ksplit) {
    constexpr bool use_a[TASK_A] = {(MASK & 0x07) != 0, (MASK & 0x38) != 0};
    constexpr bool use_b[TASK_B] = {(MASK & 0x09) != 0, (MASK & 0x12) != 0, (MASK & 0x24) != 0};
    for (int s = kpart; s < nks; s += ksplit) {
        double a[TASK_A], b[TASK_B];
#pragma unroll
        for (int i = 0; i < TASK_A; ++i) if (use_a[i]) a[i] = lds_f64(ct + a_off[i] + s * 32);
#pragma unroll
        for (int j = 0; j < TASK_B; ++j) if (use_b[j]) b[j] = lds_f64(ct + b_off[j] + s * 32);
        if (WEIGHTED) {
            const double w = lds_f64(ct + w_off + s * 32);
#pragma unroll
            for (int j = 0; j < TASK_B; ++j) if (use_b[j]) b[j] *= w;
        }
#pragma unroll
        for (int i = 0; i < TASK_A; ++i)
#pragma unroll
            for (int j = 0; j < TASK_B; ++j)
                if ((MASK >> (i * TASK_B + j)) & 1) dmma884(acc[i * TASK_B + j][0], acc[i * TASK_B + j][1], a[i], b[j]);
    }
}

template <bool WEIGHTED>
__device__ __forceinline__ void gram_ksteps_dispatch(int mask, double (&acc)[TASK_PAIRS][2], unsigned ct,
                                                     const unsigned (&a_off)[TASK_A], const unsigned (&b_off)[TASK_B],
                                                     unsigned w_off, int kpart, int nks, int ksplit) {
    switch (mask) {
#define GK(M) case M: gram_ksteps<M, WEIGHTED>(acc, ct, a_off, b_off, w_off, kpart, nks, ksplit); break;
#define GK8(B) GK(B) GK(B + 1) GK(B + 2) GK(B + 3) GK(B + 4) GK(B + 5) GK(B + 6) GK(B + 7)
        GK8(0) GK8(8) GK8(16) GK8(24) GK8(32) GK8(40) GK8(48) GK8(56)
#undef GK8
#undef GK
        default: break;
    }
}

// named barriers (bar.sync / bar.arrive with an explicit thread count): producer <-> consumer hand-over of the two
// column-tile buffers, and a consumer-only barrier for the ksplit reduction
#define BAR_FULL 1           // + buffer: producers arrive, consumers wait
#define BAR_EMPTY 3          // + buffer: consumers arrive, producers wait
#define BAR_CONS 5
__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// Warp-specialised: the first `ncons` warps are CONSUMERS (one pass-A task each, DMMA only), the remaining warps are
// PRODUCERS (TMA token tiles -> column tile).  Two column-tile buffers: the producers build chunk k + 1 while the
// consumers run the tensor cores on chunk k, so the FP64 tensor pipe never waits for term assembly.
template <bool WEIGHTED>
__global__ void __launch_bounds__(GRAM_MAX_THREADS, 1) gram_dmma_kernel(const __grid_constant__ GramParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint64_t s_bar[2];
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
    const int ncons = p.ntasks * p.ksplit, nprod = nwarps - ncons;
    const int tstride = (p.box_a * p.box_b * p.box_c + 15) & ~15;           // doubles per token tile (128-byte multiple)
    double* tiles = (double*)smem_raw;                                       // [nstages][U][box_a][box_b][box_c]
    double* ctp = tiles + (size_t)p.nstages * p.U * tstride;                 // [2][nbuild][PP_A]
    ColDesc* s_cols = (ColDesc*)(ctp + (size_t)2 * p.nbuild * PP_A);         // [nbuild]
    double* scratch = (double*)(s_cols + p.nbuild);                          // [ntasks][ksplit - 1][TASK_PAIRS * 2][32]
    load_cols_to_smem(s_cols, p.cols, p.nbuild, tstride * 8, tid, nthreads);
    if (tid == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const unsigned ct0 = smem_addr(ctp), ctb = (unsigned)(p.nbuild * PP_A * 8), cols_a = smem_addr(s_cols);

    if (warp >= ncons) {
        // =========================== producers ===========================
        const int pw_ = warp - ncons, ptid = tid - ncons * 32, pthreads = nprod * 32;
        unsigned phase[2] = {0, 0};
        int it = 0, chunk = 0;
        auto issue = [&](int ui, int st) {
            const UnitDesc u = p.g.units[ui];
            double* tile = tiles + (size_t)st * p.U * tstride;
            const int shift = u.col_lo & 1;
            if (p.use_tma) {
                if (ptid == 0) {
                    fence_proxy_async();
                    mbar_expect_tx(&s_bar[st], (unsigned)(p.U * p.box_a * p.box_b * p.box_c * 8));
                    for (int us = 0; us < p.U; ++us) {
                        double* dst = tile + (size_t)us * tstride;
                        if (p.g.ndim == 3) tma_load_tile(dst, &p.tmap[us], 3, u.col_lo - shift, u.b_lo, u.a_lo, &s_bar[st]);
                        else if (p.g.ndim == 2) tma_load_tile(dst, &p.tmap[us], 2, u.col_lo - shift, u.b_lo, 0, &s_bar[st]);
                        else tma_load_tile(dst, &p.tmap[us], 1, u.col_lo - shift, 0, 0, &s_bar[st]);
                    }
                }
            } else {
                const int wb = u.b_hi - u.b_lo, nrows = (u.a_hi - u.a_lo) * wb;
                const int total = p.U * nrows * u.ncols;
                for (int q = ptid; q < total; q += pthreads) {
                    const int cc = q % u.ncols, rr = q / u.ncols;
                    const int us = rr / nrows, row = rr - us * nrows;
                    const int ra = row / wb, rb = row - ra * wb;
                    const long long off = (long long)(u.a_lo + ra) * p.g.strideA + (long long)(u.b_lo + rb) * p.g.strideB + u.col_lo + cc;
                    tile[(size_t)us * tstride + (ra * p.box_b + rb) * p.box_c + shift + cc] = __ldg(p.tok[us] + off);
                }
            }
        };
        if (p.nstages == 2 && (int)blockIdx.x < p.g.n_units) issue(blockIdx.x, 0);
        for (int ui = blockIdx.x; ui < p.g.n_units; ui += gridDim.x, ++it) {
            const UnitDesc u = p.g.units[ui];
            const int st = (p.nstages == 2) ? (it & 1) : 0;
            if (p.nstages == 2) {
                // the other stage was last read while building the previous unit: every producer is past that
                bar_sync(BAR_CONS + 1, pthreads);
                if (ui + (int)gridDim.x < p.g.n_units) issue(ui + gridDim.x, st ^ 1);
            } else {
                bar_sync(BAR_CONS + 1, pthreads);
                issue(ui, 0);
            }
            if (p.use_tma) { mbar_wait(&s_bar[st], phase[st]); phase[st] ^= 1; }
            else bar_sync(BAR_CONS + 1, pthreads);
            // the tile address is re-defined by a volatile statement AFTER the wait: the (non-volatile, freely scheduled)
            // tile loads of build_columns depend on it and so cannot move above the arrival of the data
            unsigned tile_a = smem_addr(tiles + (size_t)st * p.U * tstride);
            asm volatile("mov.u32 %0, %0;" : "+r"(tile_a) :: "memory");
            const int wb = u.b_hi - u.b_lo;
            const float inv_wb = 1.0f / (float)wb;
            const int nrows = (u.a_hi - u.a_lo) * wb;
            const int shift = u.col_lo & 1;
            for (int q = 0; q < u.npieces; ++q) {
                const int ps = p.g.piece_start[u.piece_lo + q];
                const int pw = p.g.piece_start[u.piece_lo + q + 1] - ps;
                const int c_lo = ps - u.col_lo + shift;
                const int npts = nrows * pw;
                const float inv_pw = 1.0f / (float)pw;
                for (int c0 = 0; c0 < npts; c0 += PT_A, ++chunk) {
                    const int bf = chunk & 1;
                    int toffb[PT_A / 32];
                    bool valid[PT_A / 32];
#pragma unroll
                    for (int k = 0; k < PT_A / 32; ++k) {
                        const int pi = c0 + lane + 32 * k;
                        valid[k] = pi < npts;
                        const int r = small_div(pi, pw, inv_pw);
                        const int ra = small_div(r, wb, inv_wb);
                        toffb[k] = valid[k] ? (((ra * p.box_b + (r - ra * wb)) * p.box_c + c_lo + (pi - r * pw)) * 8) : 0;
                    }
                    bar_sync(BAR_EMPTY + bf, nthreads);                  // consumers are done with this buffer
                    build_columns<PT_A / 32>(ct0 + bf * ctb, PP_A * 8, cols_a, p.nbuild, p.gfac, tile_a, tstride * 8, toffb, valid,
                                             pw_, nprod, lane, chunk);
                    __threadfence_block();
                    bar_arrive(BAR_FULL + bf, nthreads);
                }
            }
        }
        return;
    }

    // =========================== consumers ===========================
    const int task = warp / p.ksplit, kpart = warp % p.ksplit;
    const TaskDesc td = p.tasks[task];
    unsigned a_off[TASK_A], b_off[TASK_B];
    const unsigned frag = (unsigned)(((lane >> 2) * PP_A + (lane & 3)) * 8);
#pragma unroll
    for (int i = 0; i < TASK_A; ++i) a_off[i] = frag + (unsigned)(td.ablk[i] * 8 * PP_A * 8);
#pragma unroll
    for (int j = 0; j < TASK_B; ++j) b_off[j] = frag + (unsigned)(td.bblk[j] * 8 * PP_A * 8);
    const unsigned w_off = (unsigned)((p.wcol * PP_A + (lane & 3)) * 8);
    bool on[TASK_PAIRS];
#pragma unroll
    for (int m = 0; m < TASK_PAIRS; ++m) on[m] = (td.mask >> m) & 1;
    const int* fmap = p.flushmap + ((size_t)task * TASK_PAIRS * 32 + lane) * 2;
    const int cthreads = ncons * 32;
    bar_arrive(BAR_EMPTY + 0, nthreads);                                  // both buffers start empty
    bar_arrive(BAR_EMPTY + 1, nthreads);
    int chunk = 0;
    for (int ui = blockIdx.x; ui < p.g.n_units; ui += gridDim.x) {
        const UnitDesc u = p.g.units[ui];
        const int nrows = (u.a_hi - u.a_lo) * (u.b_hi - u.b_lo);
        for (int q = 0; q < u.npieces; ++q) {
            const int pw = p.g.piece_start[u.piece_lo + q + 1] - p.g.piece_start[u.piece_lo + q];
            const int npts = nrows * pw;
            double acc[TASK_PAIRS][2];
#pragma unroll
            for (int m = 0; m < TASK_PAIRS; ++m) { acc[m][0] = 0.0; acc[m][1] = 0.0; }
            for (int c0 = 0; c0 < npts; c0 += PT_A, ++chunk) {
                const int bf = chunk & 1;
                const unsigned ct = ct0 + bf * ctb;
                bar_sync(BAR_FULL + bf, nthreads);                        // the producers have built this buffer
                const int nks = (min(PT_A, npts - c0) + 3) >> 2;
                gram_ksteps_dispatch<WEIGHTED>(td.mask, acc, ct, a_off, b_off, w_off, kpart, nks, p.ksplit);
                bar_arrive(BAR_EMPTY + bf, nthreads);
            }
            // ---- flush the run
            double* out = p.cells + ((long long)u.cell * p.g.npieces + u.piece_lo + q) * p.n_ent;
            if (p.ksplit > 1) {
                if (kpart > 0) {
                    double* sc = scratch + ((size_t)(task * (p.ksplit - 1) + kpart - 1) * TASK_PAIRS * 2) * 32 + lane;
#pragma unroll
                    for (int m = 0; m < TASK_PAIRS; ++m) { sc[(2 * m) * 32] = acc[m][0]; sc[(2 * m + 1) * 32] = acc[m][1]; }
                }
                bar_sync(BAR_CONS, cthreads);
                if (kpart == 0) {
                    for (int k = 1; k < p.ksplit; ++k) {
                        const double* sc = scratch + ((size_t)(task * (p.ksplit - 1) + k - 1) * TASK_PAIRS * 2) * 32 + lane;
#pragma unroll
                        for (int m = 0; m < TASK_PAIRS; ++m) { acc[m][0] += sc[(2 * m) * 32]; acc[m][1] += sc[(2 * m + 1) * 32]; }
                    }
                }
            }
            if (kpart == 0) {
#pragma unroll
                for (int m = 0; m < TASK_PAIRS; ++m) {
                    if (!on[m]) continue;
                    const int e0 = __ldg(fmap + m * 64), e1 = __ldg(fmap + m * 64 + 1);
                    if (e0 >= 0) out[e0] = acc[m][0];
                    if (e1 >= 0) out[e1] = acc[m][1];
                }
            }
            if (p.ksplit > 1) bar_sync(BAR_CONS, cthreads);               // scratch is reused by the next run
        }
    }
}

// ------------------------------------------------------------------------------------------
// marginals: segment row = sum of the cell records of its slab, in a fixed order (deterministic).
// grid = (n_rows, ceil(n_ent / 128)); thread <-> pair entry, so a warp reads 256 contiguous bytes of every cell
// record; four independent partial sums keep loads in flight.
// ------------------------------------------------------------------------------------------
#define MARG_SPLIT 8           // partial sums per segment row (more loads in flight; summed in split order afterwards)
struct MargParams {
    Geom g;
    const double* cells;
    double* part;            // [MARG_SPLIT][n_rows][n_ent]
    double* seg;             // [n_rows][n_ent]
    int n_ent;
};

// grid = (n_rows, ceil(n_ent / 128), MARG_SPLIT)
__global__ void __launch_bounds__(128) marginal_kernel(MargParams p) {
    const int ent = blockIdx.y * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    const MargDesc m = p.g.marg[blockIdx.x];
    const int nb = m.b_hi - m.b_lo, np = m.p_hi - m.p_lo;
    const long long inner = (long long)nb * np;
    const long long total = (long long)(m.a_hi - m.a_lo) * inner;
    const long long per = (total + MARG_SPLIT - 1) / MARG_SPLIT;
    const long long lo = (long long)blockIdx.z * per, hi = min(total, lo + per);
    const double* C = p.cells + ent;
    auto at = [&](long long q) -> double {
        const int ia = (int)(q / inner);
        const int r = (int)(q - (long long)ia * inner);
        const int ib = r / np, ip = r - ib * np;
        const long long cell = ((long long)(m.a_lo + ia) * p.g.nB + (m.b_lo + ib)) * p.g.npieces + (m.p_lo + ip);
        return __ldcs(C + cell * p.n_ent);
    };
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    long long q = lo;
    for (; q + 3 < hi; q += 4) {
        const double v0 = at(q), v1 = at(q + 1), v2 = at(q + 2), v3 = at(q + 3);
        s0 += v0; s1 += v1; s2 += v2; s3 += v3;
    }
    for (; q < hi; ++q) s0 += at(q);
    p.part[((long long)blockIdx.z * p.g.n_rows + m.row) * p.n_ent + ent] = (s0 + s1) + (s2 + s3);
}

// grid = (n_rows, ceil(n_ent / 128))
__global__ void __launch_bounds__(128) marginal_final_kernel(MargParams p) {
    const int ent = blockIdx.y * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    const int row = blockIdx.x;
    double s = 0.0;
#pragma unroll
    for (int z = 0; z < MARG_SPLIT; ++z) s += p.part[((long long)z * p.g.n_rows + row) * p.n_ent + ent];
    p.seg[(long long)row * p.n_ent + ent] = s;
}

// Grids with two leading slots (3-D): every cell belongs to one segment row of EACH axis, so the kernel above reads the
// cell table three times.  Two steps read it once: step 1, one CTA per (slot-A cell index a, 128 entries, chunk z of the
// slot-B range), walks the contiguous slab [a][b in chunk z][all pieces] and leaves  tot[a][z], byB[a][b], byP[a][z][piece]
// (byP accumulates in shared memory, one column per thread); step 2 sums those over the ranges of each segment row, in
// index order.  The chunks exist to put enough CTAs (and loads) in flight when nA x entry blocks alone is a small grid.
#define MARG_NBC 4
struct MargSlabParams {
    Geom g;
    const double* cells;
    double* part;            // [nA][MARG_NBC + nB + MARG_NBC * npieces][n_ent]
    double* seg;
    int n_ent;
};

// grid = (nA, ceil(n_ent / 128), MARG_NBC); dynamic shared memory: npieces * 128 doubles
__global__ void __launch_bounds__(128) marginal_slab_kernel(MargSlabParams p) {
    extern __shared__ double s_pp[];               // [npieces][128]
    const int a = blockIdx.x, tid = threadIdx.x, z = blockIdx.z;
    const int ent = blockIdx.y * 128 + tid;
    const int nB = p.g.nB, np = p.g.npieces;
    for (int q = 0; q < np; ++q) s_pp[q * 128 + tid] = 0.0;
    if (ent >= p.n_ent) return;
    const long long ne = p.n_ent;
    const int per = (nB + MARG_NBC - 1) / MARG_NBC;
    const int b_lo = z * per, b_hi = min(nB, b_lo + per);
    const double* C = p.cells + (long long)a * nB * np * ne + ent;
    double* out = p.part + (long long)a * (MARG_NBC + nB + MARG_NBC * np) * ne + ent;
    double tot = 0.0;
    for (int b = b_lo; b < b_hi; ++b) {
        const double* Cb = C + (long long)b * np * ne;
        double v0 = 0.0, v1 = 0.0;
        int q = 0;
        for (; q + 7 < np; q += 8) {
            double x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = __ldcs(Cb + (long long)(q + k) * ne);
#pragma unroll
            for (int k = 0; k < 8; ++k) s_pp[(q + k) * 128 + tid] += x[k];
            v0 += (x[0] + x[1]) + (x[2] + x[3]);
            v1 += (x[4] + x[5]) + (x[6] + x[7]);
        }
        for (; q < np; ++q) { const double x = __ldcs(Cb + (long long)q * ne); s_pp[q * 128 + tid] += x; v0 += x; }
        const double vb = v0 + v1;
        out[(long long)(MARG_NBC + b) * ne] = vb;
        tot += vb;
    }
    out[(long long)z * ne] = tot;
    for (int q = 0; q < np; ++q) out[(long long)(MARG_NBC + nB + z * np + q) * ne] = s_pp[q * 128 + tid];
}

// grid = (n_rows, ceil(n_ent / 128))
__global__ void __launch_bounds__(128) marginal_rows_kernel(MargSlabParams p) {
    const int ent = blockIdx.y * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    const MargDesc m = p.g.marg[blockIdx.x];
    const int nB = p.g.nB, np = p.g.npieces;
    const long long ne = p.n_ent;
    const long long rowlen = (long long)(MARG_NBC + nB + MARG_NBC * np) * ne;
    const double* P = p.part + ent;
    double s = 0.0;
    if (m.kind == 0) {
        for (int a = m.a_lo; a < m.a_hi; ++a) {
            double t = 0.0;
#pragma unroll
            for (int z = 0; z < MARG_NBC; ++z) t += P[a * rowlen + z * ne];
            s += t;
        }
    } else if (m.kind == 1) {
        for (int b = m.b_lo; b < m.b_hi; ++b) {
            double t0 = 0.0, t1 = 0.0;
            int a = 0;
            for (; a + 1 < p.g.nA; a += 2) {
                t0 += P[a * rowlen + (long long)(MARG_NBC + b) * ne];
                t1 += P[(a + 1) * rowlen + (long long)(MARG_NBC + b) * ne];
            }
            if (a < p.g.nA) t0 += P[a * rowlen + (long long)(MARG_NBC + b) * ne];
            s += t0 + t1;
        }
    } else {
        for (int q = m.p_lo; q < m.p_hi; ++q) {
            double t0 = 0.0, t1 = 0.0;
            for (int a = 0; a < p.g.nA; ++a) {
                const double* Pa = P + a * rowlen + (long long)(MARG_NBC + nB + q) * ne;
                t0 += Pa[0] + Pa[(long long)np * ne];
                t1 += Pa[(long long)2 * np * ne] + Pa[(long long)3 * np * ne];
            }
            s += t0 + t1;
        }
    }
    p.seg[(long long)m.row * p.n_ent + ent] = s;
}
static_assert(MARG_NBC == 4, "marginal_rows_kernel sums four slot-B chunks explicitly");

// window row = direct sum of the segment rows of the window; row W = every point (all rows of axis 0).
// grid = (W + 1, ceil(n_ent / 128))
struct WindowParams {
    Geom g;
    const double* seg;       // [n_rows][n_ent]
    double* win;             // [W + 1][n_ent]
    const double* total;     // weighted mode: unweighted totals [n_ent] for row W, else null
    int n_ent;
};

__global__ void __launch_bounds__(128) window_gram_kernel(WindowParams p) {
    const int ent = blockIdx.y * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    const int w = blockIdx.x;
    double s = 0.0;
    if (w == p.g.n_windows && p.total) {
        s = p.total[ent];
    } else {
        const int r0 = p.g.win_lo_row[w], r1 = p.g.win_hi_row[w];
        for (int r = r0; r < r1; ++r) s += p.seg[(long long)r * p.n_ent + ent];
    }
    p.win[(long long)w * p.n_ent + ent] = s;
}

// sum of ALL cell records (weighted mode: the unweighted total Gram of the LASSO sums, sparsity.py:54-55,73-76)
// in two fixed-order steps: grid = (parts, ceil(n_ent / 128)) -> part sums, then one CTA row sums the parts.
struct TotalParams {
    const double* cells;     // [n_cells][n_ent]
    double* part;            // [parts][n_ent]
    double* total;           // [n_ent]
    int n_cells, n_ent, parts;
};

__global__ void __launch_bounds__(128) total_part_kernel(TotalParams p) {
    const int ent = blockIdx.y * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    const long long per = ((long long)p.n_cells + p.parts - 1) / p.parts;
    const long long lo = (long long)blockIdx.x * per, hi = min((long long)p.n_cells, lo + per);
    double s0 = 0.0, s1 = 0.0;
    long long q = lo;
    for (; q + 1 < hi; q += 2) { s0 += __ldcs(p.cells + q * p.n_ent + ent); s1 += __ldcs(p.cells + (q + 1) * p.n_ent + ent); }
    if (q < hi) s0 += __ldcs(p.cells + q * p.n_ent + ent);
    p.part[(long long)blockIdx.x * p.n_ent + ent] = s0 + s1;
}

__global__ void __launch_bounds__(128) total_final_kernel(TotalParams p) {
    const int ent = blockIdx.x * 128 + threadIdx.x;
    if (ent >= p.n_ent) return;
    double s = 0.0;
    for (int q = 0; q < p.parts; ++q) s += p.part[(long long)q * p.n_ent + ent];
    p.total[ent] = s;
}

// ------------------------------------------------------------------------------------------
// solve: one CTA per eval.  calculate_weights (supplementary.py:455-469) = every window's ridge
// system solved by a GS-lane group (Gaussian elimination with partial pivoting, as LAPACK gesv);
// PhysicsInformedLasso.fit (sparsity.py:42-150) on the total Gram by thread 0 with the reference's
// exact control flow.  The equation reads the launch group's pair table through its column list.
// ------------------------------------------------------------------------------------------
struct SolveParams {
    Geom g;
    const EqDesc* eqs;
    const int* eval_list;
    const int* eval_eq;
    const int* eval_tgt;
    const double* win;       // window tables [W + 1][n_ent] per launch group (row W = all points)
    const int* pairent;      // (column, column) -> entry lookup tables per launch group
    const int* mode;         // per eval: 0 LASSO, 1 given weights + window solves, 2 given weights only (may be null)
    const double* given_wterm;   // [n_eval][TMAX] coefficient per term index (modes 1, 2)
    const double* given_b;       // [n_eval]
    // outputs (per eval)
    double* coef;            // [n_eval][TMAX]   feature order
    double* intercept;       // [n_eval]
    double* intercept_l1;    // [n_eval] offset used by the L1 fitness: intercept, or mean(target) when no
                             //          feature survives (fitness.py:135-136 'features is None')
    double* stability;       // [n_eval]
    double* sw;              // [n_eval][W][KS]
    double* wterm;           // [n_eval][TMAX]   coefficient per TERM index (0 for target / inactive)
    double* n_active;        // counts are stored as doubles: the whole solution block is one float64 all-reduce under
    double* all_zero;        // grid sharding (epde_batch_solve_part)
    double* iters;
    int part, nparts;        // this launch solves the evals at list positions part, part + nparts, ...
    int stage;               // GS = 8: keep a shared-memory copy of the equation's window Grams (dynamic smem sized for it)
};

#define SOLVE_THREADS 128
// threads per CTA of lasso_solve_kernel<GS>: systems of <= 8 unknowns (equations of <= 7 terms, the common case) get 256
// threads = 32 window systems in flight and a shared-memory copy of the window Grams of their equation
__host__ __device__ constexpr int solve_threads(int gs) { return gs == 8 ? 256 : SOLVE_THREADS; }
#define SOLVE_STAGE_DIM 9              // GS = 8: T <= 8 terms + the ones column
#define SOLVE_STAGE_DOUBLES (SOLVE_STAGE_DIM * SOLVE_STAGE_DIM)

__device__ __forceinline__ double nan_to_num(double x) {
    if (isnan(x)) return 0.0;
    if (isinf(x)) return x > 0 ? 1.7976931348623157e308 : -1.7976931348623157e308;
    return x;
}

template <int GS>
__global__ void __launch_bounds__(solve_threads(GS)) lasso_solve_kernel(SolveParams p) {
    constexpr int NT = solve_threads(GS);
    constexpr int NG = NT / GS;
    __shared__ double Gt[KS * KS];             // total Gram of all points, symmetric K x K
    __shared__ double M[NG][GS][GS + 1];       // augmented systems
    __shared__ int s_ent[KS * KS];             // (term | ones, term | ones) -> entry of the group's pair table
    __shared__ int cols[KS];                   // columns of the current system: active terms, then T (ones)
    __shared__ double s_coef[TMAX], s_cv[TMAX], s_mu[KS], s_cvraw[KS];
    __shared__ int s_n, s_state, s_iter, s_sweeps;
    __shared__ double s_b;
    extern __shared__ double s_sw[];           // [W][KS]

    const int tid = threadIdx.x;
    const int ev = p.eval_list[blockIdx.x * p.nparts + p.part];
    const int e = p.eval_eq[ev];
    const int tgt = p.eval_tgt[ev];
    const EqDesc& eq = p.eqs[e];
    const int T = eq.T, F = T - 1;
    const int NE = eq.n_ent;
    const int W = p.g.n_windows;
    const double* P = p.win + eq.w_off;
    const int* pe = p.pairent + eq.pe_off;

    for (int q = tid; q < (T + 1) * (T + 1); q += NT) {
        const int i = q / (T + 1), j = q % (T + 1);
        const int en = pe[(int)eq.col[i] * eq.cp + (int)eq.col[j]];
        s_ent[i * KS + j] = en;
        Gt[i * KS + j] = P[(long long)W * NE + en];
    }
    if (tid == 0) {
        int n = 0;
        for (int t = 0; t < T; ++t)
            if (t != tgt) cols[n++] = t;
        cols[n] = T;
        s_n = n + 1;
        s_state = 0;      // 0: first fit, 1: refit after a zeroed coefficient, 2: done
        s_iter = 0;
        s_sweeps = 0;
        for (int j = 0; j < TMAX; ++j) { s_coef[j] = 0.0; s_cv[j] = 0.0; }
    }
    __syncthreads();

    // GS = 8: the (T + 1) x (T + 1) Gram of every window, copied once (every refit solves all W systems again)
    double* s_win = s_sw + (size_t)W * KS;     // [W][9][9]
    const bool staged = GS == 8 && p.stage;
    if (staged) {
        for (int q = tid; q < W * SOLVE_STAGE_DOUBLES; q += NT) {
            const int w = q / SOLVE_STAGE_DOUBLES, r = q - w * SOLVE_STAGE_DOUBLES;
            const int i = r / SOLVE_STAGE_DIM, j = r - i * SOLVE_STAGE_DIM;
            s_win[q] = (i <= T && j <= T) ? P[(long long)w * NE + s_ent[i * KS + j]] : 0.0;
        }
        __syncthreads();
    }
    const int grp = tid / GS, lr = tid % GS;
    // feature j (0..F-1) <-> term index
    auto feat_term = [&](int j) { return j < tgt ? j : j + 1; };
    const int mode = p.mode ? p.mode[ev] : 0;

    while (mode != 2) {
        const int n = s_n;                     // system size (active features + ones)
        // ---------------- calculate_weights on the active set ----------------
        const int rounds = (W + NG - 1) / NG;
        for (int it = 0; it < rounds; ++it) {
            int w = it * NG + grp;
            const bool valid = w < W;
            if (!valid) w = W - 1;
            const double* Phi = P + (long long)w * NE;
            double (*Mg)[GS + 1] = M[grp];
            if (lr < n) {
                const int i = cols[lr];
                if (staged) {
                    const double* Sw = s_win + (size_t)w * SOLVE_STAGE_DOUBLES + i * SOLVE_STAGE_DIM;
                    for (int c = 0; c < n; ++c) {
                        double v = Sw[cols[c]];
                        if (c == lr) v += RIDGE;
                        Mg[lr][c] = v;
                    }
                    Mg[lr][n] = Sw[tgt];
                } else {
                    for (int c = 0; c < n; ++c) {
                        double v = Phi[s_ent[i * KS + cols[c]]];
                        if (c == lr) v += RIDGE;
                        Mg[lr][c] = v;
                    }
                    Mg[lr][n] = Phi[s_ent[i * KS + tgt]];
                }
            }
            __syncwarp();
            for (int k = 0; k < n; ++k) {
                // partial pivoting: first row of maximal |a_rk|, r >= k
                double val = (lr >= k && lr < n) ? fabs(Mg[lr][k]) : -1.0;
                int idx = lr;
#pragma unroll
                for (int off = GS / 2; off > 0; off >>= 1) {
                    const double ov = __shfl_xor_sync(0xffffffffu, val, off, GS);
                    const int oi = __shfl_xor_sync(0xffffffffu, idx, off, GS);
                    if (ov > val || (ov == val && oi < idx)) { val = ov; idx = oi; }
                }
                if (idx != k && lr >= k && lr < n) {           // lane <-> column; lane k also swaps the right-hand side
                    const double t0 = Mg[k][lr];
                    Mg[k][lr] = Mg[idx][lr];
                    Mg[idx][lr] = t0;
                    if (lr == k) { const double t1 = Mg[k][n]; Mg[k][n] = Mg[idx][n]; Mg[idx][n] = t1; }
                }
                __syncwarp();
                if (lr > k && lr < n) {
                    const double f = Mg[lr][k] / Mg[k][k];
                    for (int c = k + 1; c <= n; ++c) Mg[lr][c] = fma(-f, Mg[k][c], Mg[lr][c]);
                }
                __syncwarp();
            }
            for (int k = n - 1; k >= 0; --k) {
                const double xk = Mg[k][n] / Mg[k][k];
                __syncwarp();
                if (lr < k) Mg[lr][n] = fma(-Mg[lr][k], xk, Mg[lr][n]);
                if (lr == k && valid) s_sw[w * KS + k] = xk;
                __syncwarp();
            }
        }
        __syncthreads();
        // ---------------- column statistics: mean, var(ddof=1), CV (sparsity.py:29-40) ----------
        // one warp per column, lanes stride over the windows, butterfly sums (fixed order)
        for (int c = tid >> 5; c < n; c += NT / 32) {
            const int lane = tid & 31;
            double s = 0.0;
            for (int w = lane; w < W; w += 32) s += s_sw[w * KS + c];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
            const double mu = s / W;
            double q = 0.0;
            for (int w = lane; w < W; w += 32) { const double dlt = s_sw[w * KS + c] - mu; q = fma(dlt, dlt, q); }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) q += __shfl_xor_sync(0xffffffffu, q, off);
            const double var = q / (W - 1);     // std(ddof=1)**2 ; W == 1 -> nan like numpy
            double cv = var / (mu * mu);
            if (mu == 0.0) cv = 0.0;
            if (lane == 0) { s_mu[c] = mu; s_cvraw[c] = cv; }
        }
        __syncthreads();
        // ---------------- coordinate descent (thread 0) ----------------
        if (tid == 0 && mode == 1) s_state = 2;   // in-place fitness: window weights only (fitness.py:155-159)
        if (tid == 0 && mode == 0) {
            // scatter the refit into full-length feature vectors (sparsity.py:49-52, 87-96)
            {
                int a = 0;
                for (int j = 0; j < F; ++j) {
                    const bool act = (a < n - 1) && (cols[a] == feat_term(j));
                    if (act) { s_coef[j] = s_mu[a]; s_cv[j] = nan_to_num(s_cvraw[a]); ++a; }
                    else { s_coef[j] = 0.0; s_cv[j] = 0.0; }
                }
                s_b = s_mu[n - 1];
            }
            const double yy = Gt[tgt * KS + tgt], sy = Gt[tgt * KS + T], Npts = Gt[T * KS + T];
            int iteration = s_iter;
            bool refit = false;
            bool first_pass_after_refit = (s_state == 1);
            // after a refit the reference finishes the interrupted sweep's bookkeeping:
            // max_change = inf fails criterion 1, then iteration += 1  (sparsity.py:96-150)
            if (first_pass_after_refit) iteration = 1;
            // Within one fit the intercept, the CVs and the SET of non-zero coefficients are fixed (a coefficient that
            // reaches zero ends the fit at once), so everything that depends only on them is computed once per fit:
            // the visiting order argsort(cv)[::-1] (descending CV, ties -> higher index first) restricted to the
            // non-zero coefficients, y_sq_sum and the thresholds.  Skipped zero coefficients contribute exact zeros
            // to every sum of the reference, so the values are unchanged.
            int act[TMAX], na = 0;
            for (int j = 0; j < F; ++j)
                if (s_coef[j] != 0.0) act[na++] = j;
            for (int a = 1; a < na; ++a) {
                const int oj = act[a];
                const double cvj = s_cv[oj];
                int b2 = a - 1;
                while (b2 >= 0 && (s_cv[act[b2]] < cvj || (s_cv[act[b2]] == cvj && act[b2] < oj))) {
                    act[b2 + 1] = act[b2];
                    --b2;
                }
                act[b2 + 1] = oj;
            }
            const double y_sq_sum = yy - 2.0 * s_b * sy + Npts * s_b * s_b;
            auto xt_res = [&](int j) {
                const int tj = feat_term(j);
                double s = Gt[tj * KS + tgt];
                // ascending feature index, as the reference's dot product runs
                for (int k2 = 0; k2 < F; ++k2) {
                    const double ck = s_coef[k2];
                    if (ck != 0.0) s = fma(-Gt[tj * KS + feat_term(k2)], ck, s);
                }
                return fma(-s_b, Gt[tj * KS + T], s);
            };
            while (iteration < LASSO_MAX_ITER && na > 0) {
                ++s_sweeps;
                double max_change = 0.0;
                for (int q = 0; q < na; ++q) {
                    const int j = act[q];
                    const double old_coef = s_coef[j];
                    const int tj = feat_term(j);
                    const double norm_sq = Gt[tj * KS + tj];
                    const double rho = xt_res(j) + old_coef * norm_sq;
                    const double thr = s_cv[j] * y_sq_sum;
                    const double mag = fmax(fabs(rho) - thr, 0.0);
                    const double sgn = (rho > 0.0) ? 1.0 : ((rho < 0.0) ? -1.0 : 0.0);
                    const double new_coef = (sgn * mag) / norm_sq;
                    s_coef[j] = new_coef;
                    if (new_coef == 0.0) { refit = true; break; }
                    const double change = fabs(new_coef - old_coef);
                    if (change > max_change) max_change = change;
                }
                if (refit) break;
                double max_abs = 0.0;
                for (int q = 0; q < na; ++q) max_abs = fmax(max_abs, fabs(s_coef[act[q]]));
                if (max_change <= LASSO_TOL * max_abs) {
                    double dual_norm = 0.0;
                    for (int j = 0; j < F; ++j)
                        if (s_cv[j] * y_sq_sum > 0.0)
                            dual_norm = fmax(dual_norm, fabs(xt_res(j)) / s_cv[j] * y_sq_sum);
                    double cGc = 0.0, cXy = 0.0, cXs = 0.0, pen = 0.0;
                    for (int j = 0; j < F; ++j) {
                        const int tj = feat_term(j);
                        double row = 0.0;
                        for (int k2 = 0; k2 < F; ++k2) row = fma(Gt[tj * KS + feat_term(k2)], s_coef[k2], row);
                        cGc = fma(s_coef[j], row, cGc);
                        cXy = fma(s_coef[j], Gt[tj * KS + tgt], cXy);
                        cXs = fma(s_coef[j], Gt[tj * KS + T], cXs);
                        pen += s_cv[j] * y_sq_sum * fabs(s_coef[j]);
                    }
                    const double b = s_b;
                    const double rr = yy - 2.0 * cXy - 2.0 * b * sy + cGc + 2.0 * b * cXs + Npts * b * b;
                    const double sum_r = sy - cXs - Npts * b;
                    const double y_r = yy - cXy - b * sy;
                    const double yb_r = y_r - b * sum_r;
                    const double alpha = dual_norm > 1.0 ? 1.0 / dual_norm : 1.0;
                    const double primal = 0.5 * rr + pen;
                    const double dual = 0.5 * y_sq_sum - 0.5 * (y_sq_sum - 2.0 * alpha * yb_r + alpha * alpha * rr);
                    if (primal - dual <= LASSO_TOL * (y_sq_sum / Npts)) break;
                }
                ++iteration;
            }
            if (refit) {
                int nn = 0;
                for (int j = 0; j < F; ++j)
                    if (s_coef[j] != 0.0) cols[nn++] = feat_term(j);
                cols[nn] = T;
                s_n = nn + 1;
                s_state = 1;
                s_iter = 0;
            } else {
                s_state = 2;
            }
        }
        __syncthreads();
        if (s_state == 2) break;
    }

    // ---------------- outputs ----------------
    if (tid == 0 && mode != 0) {
        for (int j = 0; j < F; ++j) s_coef[j] = p.given_wterm[(long long)ev * TMAX + feat_term(j)];
        s_b = p.given_b[ev];
    }
    __syncthreads();
    const int n = (mode == 2) ? 0 : s_n;
    for (int q = tid; q < W * KS; q += NT) {
        const int c = q % KS;
        p.sw[(long long)ev * W * KS + q] = (c < n) ? s_sw[q] : 0.0;
    }
    if (tid == 0) {
        int nact = 0;
        for (int j = 0; j < TMAX; ++j) {
            const double cj = (j < F) ? s_coef[j] : 0.0;
            p.coef[(long long)ev * TMAX + j] = cj;
            if (cj != 0.0) ++nact;
        }
        for (int t = 0; t < TMAX; ++t) p.wterm[(long long)ev * TMAX + t] = 0.0;
        for (int j = 0; j < F; ++j) p.wterm[(long long)ev * TMAX + feat_term(j)] = s_coef[j];
        p.intercept[ev] = s_b;
        p.intercept_l1[ev] = (nact == 0) ? Gt[tgt * KS + T] / Gt[T * KS + T] : s_b;
        p.n_active[ev] = (double)nact;
        p.all_zero[ev] = (nact == 0) ? 1.0 : 0.0;
        p.iters[ev] = (double)s_sweeps;
        // stability = sum(cv[:-1]) / ndim over the columns of the cached weights (fitness.py:160-169)
        double st = 0.0;
        for (int a = 0; a < n - 1; ++a) st += s_cvraw[a];
        p.stability[ev] = (mode == 2) ? nan("") : st / p.g.ndim;
    }
}

// ------------------------------------------------------------------------------------------
// pass B: residual sums per eval: sum|r|, sum|t|, sum r2^2 with r = t - (F w + b1), r2 = t - (F w + b)
// (fitness.py:135-142 and 70-83).  Persistent CTAs (one per SM) walk over chunks of PT_B consecutive points
// (token rows of the chunk fetched with one bulk async copy per token, two stages).  Phase 1 builds the column tile
// of the launch group's dictionary once per chunk; phase 2: warps take jobs from a shared counter:
//   octet job   8 consecutive evals of the group's list (sorted by target column): lane <-> 4 points per 128-point
//               sweep; a residual is the sparse dot of the eval's target and surviving columns read from the column
//               tile (the target column stays in registers while consecutive evals share it); the 8 per-eval sums are
//               reduced over the warp together (reduce-scatter butterfly: 9 shuffle steps instead of 40);
//   target job  sum|t| of one distinct target column.
// Chunk sums go to a per-CTA shared table in chunk order, written once per CTA and reduced over CTAs in CTA order by
// residual_final_kernel: fixed summation order, no atomics on data.
// ------------------------------------------------------------------------------------------
#define OCT 8                // evals per octet job

struct EvalJob {             // 32 bytes, built on the device from the solve outputs
    int tcol, nnz, pad0, pad1;   // target column (byte offset of its row in the column tile), surviving features
    double b1;               // offset of the L1 residual
    double delta;            // r2 = r + delta (b1 - intercept: non-zero only when no feature survives)
};

struct ResidProgParams {
    const EqDesc* eqs;
    const int* eval_eq;
    const int* eval_tgt;
    const int* slot_ev;      // [n_slots] eval id of every sum slot (sorted by target column)
    const double* wterm;     // [n_eval][TMAX]
    const double* intercept;
    const double* intercept_l1;
    EvalJob* jobs;           // [n_slots rounded up to OCT]
    int* job_col;            // [n_slots][TMAX] byte offsets of the surviving columns' rows in the column tile
    double* job_w;           // [n_slots][TMAX]
    int n_slots;
};

__global__ void residual_prog_kernel(ResidProgParams p) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_pad = (p.n_slots + OCT - 1) / OCT * OCT;
    if (s >= n_pad) return;
    EvalJob j;
    j.tcol = 0; j.nnz = 0; j.pad0 = j.pad1 = 0; j.b1 = 0.0; j.delta = 0.0;
    if (s < p.n_slots) {
        const int ev = p.slot_ev[s];
        const EqDesc& eq = p.eqs[p.eval_eq[ev]];
        const int tgt = p.eval_tgt[ev];
        int nnz = 0;
        for (int t = 0; t < eq.T; ++t) {
            const double w = p.wterm[(long long)ev * TMAX + t];
            if (t != tgt && w != 0.0) {
                p.job_col[(long long)s * TMAX + nnz] = (int)eq.col[t] * (PP_B * 8);
                p.job_w[(long long)s * TMAX + nnz] = w;
                ++nnz;
            }
        }
        j.tcol = (int)eq.col[tgt] * (PP_B * 8); j.nnz = nnz;
        j.b1 = p.intercept_l1[ev];
        j.delta = p.intercept_l1[ev] - p.intercept[ev];
    }
    p.jobs[s] = j;
}

// Pass B walks the evals of a launch group sorted by (target column, first surviving column), so that consecutive
// evals of an octet reuse the registers that hold those two columns, and reads their descriptors from a COMPACT
// table (32-byte headers in that order + CSR lists of (column, weight) entries) small enough to sit in shared memory.
// The surviving sets only exist after the solve, so ordering and packing run on the device: one CTA, bitonic sort of
// 64-bit keys (column, column, slot), then a prefix sum of the entry counts.  More than SORT_MAX evals: slot order.
#define SORT_MAX 8192
struct EvalHdr {             // 32 bytes
    int tcol;                // byte offset of the target column's row in the column tile
    int nnz;                 // surviving features
    int off;                 // first entry in the CSR arrays
    int slot;                // sum slot (-1: padding)
    double b1, delta;
};

struct ResidPackParams {
    const EvalJob* jobs;     // [n_slots] in slot order
    const int* job_col;      // [n_slots][TMAX]
    const double* job_w;     // [n_slots][TMAX]
    EvalHdr* hdr;            // [n_slots rounded up to OCT] in walking order
    int* ent_col;            // [sum nnz]
    double* ent_w;           // [sum nnz]
    int* meta;               // [0] = sum nnz, [1] = number of columns in blist
    int* blist;              // [nbuild] columns of the dictionary that some eval reads (targets, surviving terms, weights)
    int nbuild, wcol;
    int n_slots;
};

__global__ void __launch_bounds__(1024) residual_pack_kernel(ResidPackParams p) {
    extern __shared__ unsigned long long s_key[];          // n2 keys, then n2 ints (scan)
    // columns pass B has to build: after the LASSO most columns of the dictionary survive in no eval
    __shared__ int s_need[CMAX + 8];
    for (int q = threadIdx.x; q < p.nbuild; q += blockDim.x) s_need[q] = 0;
    __syncthreads();
    for (int q = threadIdx.x; q < p.n_slots; q += blockDim.x) {
        const EvalJob j = p.jobs[q];
        s_need[j.tcol / (PP_B * 8)] = 1;
        for (int z = 0; z < j.nnz; ++z) s_need[p.job_col[(long long)q * TMAX + z] / (PP_B * 8)] = 1;
    }
    if (threadIdx.x == 0 && p.wcol >= 0) s_need[p.wcol] = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        int nb = 0;
        for (int c = 0; c < p.nbuild; ++c)
            if (s_need[c]) p.blist[nb++] = c;
        p.meta[1] = nb;
    }
    const int n_pad = (p.n_slots + OCT - 1) / OCT * OCT;
    int n2 = 1;
    while (n2 < p.n_slots) n2 <<= 1;
    if (n2 > SORT_MAX) {                                    // too many evals for the shared-memory sort: slot order, serial scan
        if (threadIdx.x == 0) {
            int off = 0;
            for (int q = 0; q < n_pad; ++q) {
                EvalHdr h;
                h.tcol = 0; h.nnz = 0; h.off = off; h.slot = -1; h.b1 = 0.0; h.delta = 0.0;
                if (q < p.n_slots) {
                    const EvalJob j = p.jobs[q];
                    h.tcol = j.tcol; h.nnz = j.nnz; h.slot = q; h.b1 = j.b1; h.delta = j.delta;
                    for (int z = 0; z < j.nnz; ++z) { p.ent_col[off + z] = p.job_col[(long long)q * TMAX + z]; p.ent_w[off + z] = p.job_w[(long long)q * TMAX + z]; }
                    off += j.nnz;
                }
                p.hdr[q] = h;
            }
            p.meta[0] = off;
        }
        return;
    }
    int* s_scan = (int*)(s_key + n2);
    for (int q = threadIdx.x; q < n2; q += blockDim.x) {
        unsigned long long k = ~0ull;
        if (q < p.n_slots) {
            const EvalJob j = p.jobs[q];
            const unsigned long long c0 = j.nnz > 0 ? (unsigned long long)(p.job_col[(long long)q * TMAX] / (PP_B * 8)) : 0xffffull;
            k = ((unsigned long long)(j.tcol / (PP_B * 8)) << 44) | (c0 << 24) | (unsigned long long)q;
        }
        s_key[q] = k;
    }
    __syncthreads();
    for (int k = 2; k <= n2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int q = threadIdx.x; q < n2; q += blockDim.x) {
                const int x = q ^ j;
                if (x > q) {
                    const unsigned long long a = s_key[q], b = s_key[x];
                    const bool up = (q & k) == 0;
                    if ((a > b) == up) { s_key[q] = b; s_key[x] = a; }
                }
            }
            __syncthreads();
        }
    // inclusive scan of the entry counts in walking order (Hillis-Steele, n2 <= 8192)
    for (int q = threadIdx.x; q < n2; q += blockDim.x)
        s_scan[q] = q < p.n_slots ? p.jobs[(int)(s_key[q] & 0xffffffull)].nnz : 0;
    __syncthreads();
    for (int d = 1; d < n2; d <<= 1) {
        int add[SORT_MAX / 1024];
        int m = 0;
        for (int q = threadIdx.x; q < n2; q += blockDim.x) add[m++] = q >= d ? s_scan[q - d] : 0;
        __syncthreads();
        m = 0;
        for (int q = threadIdx.x; q < n2; q += blockDim.x) s_scan[q] += add[m++];
        __syncthreads();
    }
    for (int q = threadIdx.x; q < n_pad; q += blockDim.x) {
        EvalHdr h;
        h.tcol = 0; h.nnz = 0; h.off = 0; h.slot = -1; h.b1 = 0.0; h.delta = 0.0;
        if (q < p.n_slots) {
            const int slot = (int)(s_key[q] & 0xffffffull);
            const EvalJob j = p.jobs[slot];
            h.tcol = j.tcol; h.nnz = j.nnz; h.off = s_scan[q] - j.nnz; h.slot = slot; h.b1 = j.b1; h.delta = j.delta;
            for (int z = 0; z < j.nnz; ++z) {
                p.ent_col[h.off + z] = p.job_col[(long long)slot * TMAX + z];
                p.ent_w[h.off + z] = p.job_w[(long long)slot * TMAX + z];
            }
        }
        p.hdr[q] = h;
    }
    if (threadIdx.x == 0) p.meta[0] = p.n_slots > 0 ? s_scan[p.n_slots - 1] : 0;
}

struct ResidParams {
    long long N;
    int U;
    const double* tok[MAXU + 1];
    const ColDesc* cols;
    const FactorDesc* gfac;
    int nbuild;              // columns of the tile (cp)
    const EvalHdr* hdr;      // [n_slots rounded up to OCT] in walking order (residual_pack_kernel)
    const int* ent_col;      // CSR entries: byte offsets of the surviving columns' rows in the column tile
    const double* ent_w;
    const int* meta;         // [0] = number of entries, [1] = number of columns to build
    const int* blist;        // columns to build (residual_pack_kernel)
    int wcol;                // weighted mode: row of the column tile that holds the sample weights g (L2Fitness multiplies
                             // its residual by g, fitness.py:81-82), else -1
    int desc_smem;           // bytes of shared memory available for a copy of the descriptors (0: read them from global)
    int n_slots;             // evals of the group = sum slots
    const int* tcols;        // distinct target columns
    int n_tcols;
    double* partial;         // [gridDim.x][2 * n_slots + n_tcols]
    int n_chunks;
    int nstages;
    int aligned;             // token vectors may be fetched with bulk copies
};

#define RES_NPL (PT_B / 32)

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// sums of 8 per-lane values over the warp, all at once: after the three exchange steps lane l holds value
// ((l >> 4) & 1) * 4 + ((l >> 3) & 1) * 2 + ((l >> 2) & 1) summed over the 8 lanes that share its low two bits; two more
// butterfly steps finish it.  Returns that value's warp total (the same in the 4 lanes of a group).
__device__ __forceinline__ double warp_sum8(const double (&a)[OCT], int lane) {
    double b[4], c[2];
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double keep = h16 ? a[i + 4] : a[i], give = h16 ? a[i] : a[i + 4];
        b[i] = keep + __shfl_xor_sync(0xffffffffu, give, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const double keep = h8 ? b[i + 2] : b[i], give = h8 ? b[i] : b[i + 2];
        c[i] = keep + __shfl_xor_sync(0xffffffffu, give, 8);
    }
    const double keep = h4 ? c[1] : c[0], give = h4 ? c[0] : c[1];
    double d = keep + __shfl_xor_sync(0xffffffffu, give, 4);
    d += __shfl_xor_sync(0xffffffffu, d, 2);
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    return d;
}

__device__ __forceinline__ void resid_stage_issue(const ResidParams& p, long long chunk, double* tile, uint64_t* bar,
                                                  int tid, int nthreads) {
    const long long lo = chunk * PT_B;
    const int npts = (int)min((long long)PT_B, p.N - lo);
    if (p.aligned) {
        const unsigned bytes = (unsigned)(((npts + 1) & ~1) * 8);
        if (tid < 32) {
            if (tid == 0) mbar_expect_tx(bar, bytes * (unsigned)p.U);
            __syncwarp();
            if (tid < p.U) bulk_g2s(tile + (size_t)tid * PT_B, p.tok[tid] + lo, bytes, bar);
        }
    } else {
        for (int q = tid; q < p.U * npts; q += nthreads) {
            const int us = q / npts, i = q - us * npts;
            tile[(size_t)us * PT_B + i] = __ldg(p.tok[us] + lo + i);
        }
    }
}

template <bool L2>
__global__ void __launch_bounds__(RES_WARPS * 32, 1) residual_cols_kernel(ResidParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint64_t s_bar[2];
    __shared__ int s_next;
    __shared__ int s_blist[CMAX + 8];
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
    const int n_blist = p.meta[1];
    for (int q = tid; q < n_blist; q += nthreads) s_blist[q] = p.blist[q];
    double* tiles = (double*)smem_raw;                                        // [nstages][U][PT_B]
    double* ctp = tiles + (size_t)p.nstages * p.U * PT_B;                     // [nbuild][PP_B]
    ColDesc* s_cols = (ColDesc*)(ctp + (size_t)p.nbuild * PP_B);              // [nbuild]
    double* s_sum = (double*)(s_cols + p.nbuild);                             // [2 n_slots + n_tcols]
    const int nsum = 2 * p.n_slots + p.n_tcols;
    load_cols_to_smem(s_cols, p.cols, p.nbuild, PT_B * 8, tid, nthreads);
    for (int q = tid; q < nsum; q += nthreads) s_sum[q] = 0.0;
    // descriptors of the evals: a shared-memory copy when it fits (generic pointers: the loops below do not care)
    const int n_oct = (p.n_slots + OCT - 1) / OCT;
    const EvalHdr* hdr = p.hdr;
    const int* ent_col = p.ent_col;
    const double* ent_w = p.ent_w;
    {
        const int n_ent = p.meta[0];
        const size_t need = sizeof(EvalHdr) * (size_t)n_oct * OCT + (sizeof(double) + sizeof(int)) * (size_t)n_ent + 16;
        if (need <= (size_t)p.desc_smem) {
            EvalHdr* sh = (EvalHdr*)(s_sum + ((nsum + 1) & ~1));
            double* sw = (double*)(sh + (size_t)n_oct * OCT);
            int* sc = (int*)(sw + n_ent);
            for (int q = tid; q < n_oct * OCT * 2; q += nthreads) ((double2*)sh)[q] = ((const double2*)p.hdr)[q];
            for (int q = tid; q < n_ent; q += nthreads) { sw[q] = p.ent_w[q]; sc[q] = p.ent_col[q]; }
            hdr = sh; ent_col = sc; ent_w = sw;
        }
    }
    if (tid == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const unsigned ct = smem_addr(ctp), cols_a = smem_addr(s_cols);

    const bool async = p.aligned != 0;
    unsigned phase[2] = {0, 0};
    int it = 0;
    if (p.nstages == 2 && (long long)blockIdx.x < p.n_chunks) resid_stage_issue(p, blockIdx.x, tiles, &s_bar[0], tid, nthreads);
    const int njobs_all = n_oct + p.n_tcols;
    for (long long ci = blockIdx.x; ci < p.n_chunks; ci += gridDim.x, ++it) {
        const int st = (p.nstages == 2) ? (it & 1) : 0;
        double* tile = tiles + (size_t)st * p.U * PT_B;
        if (p.nstages == 2) {
            const long long cn = ci + gridDim.x;
            if (cn < p.n_chunks) {
                if (async) fence_proxy_async();
                resid_stage_issue(p, cn, tiles + (size_t)(st ^ 1) * p.U * PT_B, &s_bar[st ^ 1], tid, nthreads);
            }
        } else {
            if (async) fence_proxy_async();
            resid_stage_issue(p, ci, tile, &s_bar[0], tid, nthreads);
        }
        if (async) { mbar_wait(&s_bar[st], phase[st]); phase[st] ^= 1; }
        else __syncthreads();

        const int npts = (int)min((long long)PT_B, p.N - ci * PT_B);
        {
            int toffb[RES_NPL];
            bool valid[RES_NPL];
#pragma unroll
            for (int k = 0; k < RES_NPL; ++k) { valid[k] = lane + 32 * k < npts; toffb[k] = valid[k] ? (lane + 32 * k) * 8 : 0; }
            unsigned tile_a = smem_addr(tile);
            asm volatile("mov.u32 %0, %0;" : "+r"(tile_a) :: "memory");     // defined after the wait: see gram_dmma_kernel
            build_columns<RES_NPL, true>(ct, PP_B * 8, cols_a, n_blist, p.gfac, tile_a, PT_B * 8, toffb, valid, warp, nwarps, lane, (int)ci,
                                         smem_addr(s_blist));
        }
        if (tid == 0) s_next = 0;
        __syncthreads();

        const unsigned cl = ct + lane * 8;
        // the next job number is drawn while the current job runs (lane 0 holds it; broadcast at the top of the loop)
        int jj_next = 0;
        if (lane == 0) jj_next = atomicAdd(&s_next, 1);
        for (;;) {
            const int jj = __shfl_sync(0xffffffffu, jj_next, 0);
            if (jj >= njobs_all) break;
            if (lane == 0) jj_next = atomicAdd(&s_next, 1);
            if (jj < n_oct) {
                // ---------------- octet job: evals [8 jj, 8 jj + 8) of the list (padding evals: nnz 0, sums discarded)
                // descriptors are fetched up front, spread over the lanes (lane v: header of eval v; lane z: entry z of
                // each eval), and broadcast by shuffles inside the loops: no global load on the critical path
                const EvalHdr mine = hdr[(size_t)jj * OCT + (lane & 7)];
                const int my_slot = mine.slot;
                const bool full = npts == PT_B;
                double a1[OCT], a2[OCT];
                int cur_t = -1, cur_c = -1;
                double t[RES_NPL], c1[RES_NPL];
#pragma unroll
                for (int v = 0; v < OCT; ++v) {
                    // every shuffle sits at the top level of the unrolled loop (none inside a conditional)
                    const int nz = __shfl_sync(0xffffffffu, mine.nnz, v);
                    const int tco = __shfl_sync(0xffffffffu, mine.tcol, v);
                    const double b1 = __shfl_sync(0xffffffffu, mine.b1, v);
                    // entry z of the eval in lane z (one coalesced fetch), broadcast by shuffle when used
                    const int eo = __shfl_sync(0xffffffffu, mine.off, v) + lane;
                    const int e_col = (lane < nz) ? ent_col[eo] : 0;
                    const double e_w = (lane < nz) ? ent_w[eo] : 0.0;
                    const int co0 = __shfl_sync(0xffffffffu, e_col, 0), co1 = __shfl_sync(0xffffffffu, e_col, 1);
                    const double w0 = __shfl_sync(0xffffffffu, e_w, 0), w1 = __shfl_sync(0xffffffffu, e_w, 1);
                    if (tco != cur_t) {
                        cur_t = tco;
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) t[k] = lds_f64_e(cl + tco + k * 256, jj);
                    }
                    double r[RES_NPL];
#pragma unroll
                    for (int k = 0; k < RES_NPL; ++k) r[k] = t[k] - b1;
                    if (nz > 0) {
                        // first surviving column: kept in registers while consecutive evals use the same one
                        if (co0 != cur_c) {
                            cur_c = co0;
#pragma unroll
                            for (int k = 0; k < RES_NPL; ++k) c1[k] = lds_f64_e(cl + co0 + k * 256, jj);
                        }
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) r[k] = fma(-w0, c1[k], r[k]);
                    }
                    if (nz > 1) {
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) r[k] = fma(-w1, lds_f64_e(cl + co1 + k * 256, jj), r[k]);
                    }
                    if (nz > 2) {
#pragma unroll 1
                        for (int z = 2; z < nz; ++z) {
                            const unsigned ca = cl + (unsigned)__shfl_sync(0xffffffffu, e_col, z);
                            const double w = __shfl_sync(0xffffffffu, e_w, z);
#pragma unroll
                            for (int k = 0; k < RES_NPL; ++k) r[k] = fma(-w, lds_f64_e(ca + k * 256, jj), r[k]);
                        }
                    }
                    double s1 = 0.0, s2 = 0.0;
                    if (L2) {
                        const double dl = __shfl_sync(0xffffffffu, mine.delta, v);
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) {
                            double r2 = r[k] + dl;
                            if (p.wcol >= 0) r2 *= lds_f64_e(cl + p.wcol * (PP_B * 8) + k * 256, jj);
                            r[k] = (full || lane + 32 * k < npts) ? r[k] : 0.0;
                            s2 = (full || lane + 32 * k < npts) ? fma(r2, r2, s2) : s2;
                        }
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) s1 += fabs(r[k]);
                    } else if (full) {
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) s1 += fabs(r[k]);
                    } else {
#pragma unroll
                        for (int k = 0; k < RES_NPL; ++k) if (lane + 32 * k < npts) s1 += fabs(r[k]);
                    }
                    a1[v] = s1; a2[v] = s2;
                }
                const int vi = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
                const int slot = __shfl_sync(0xffffffffu, my_slot, vi);
                const double t1 = warp_sum8(a1, lane);
                if ((lane & 3) == 0 && slot >= 0) s_sum[slot] += t1;
                if (L2) {
                    const double t2 = warp_sum8(a2, lane);
                    if ((lane & 3) == 0 && slot >= 0) s_sum[p.n_slots + slot] += t2;
                }
            } else {
                // ---------------- target job: sum|t| of one target column (padding points hold 0)
                const int tc = jj - n_oct;
                const int c = p.tcols[tc];
                double a = 0.0;
#pragma unroll
                for (int k = 0; k < RES_NPL; ++k) a += fabs(lds_f64(cl + c * (PP_B * 8) + k * 256));
                a = warp_sum(a);
                if (lane == 0) s_sum[2 * p.n_slots + tc] += a;
            }
        }
        __syncthreads();
    }
    double* out = p.partial + (size_t)blockIdx.x * nsum;
    for (int q = tid; q < nsum; q += nthreads) out[q] = s_sum[q];
}

struct ResidFinalParams {
    const int* slot_ev;      // [n_slots]
    const int* slot_tci;     // [n_slots] index of the eval's target column in the target-column list
    const double* partial;   // [n_cta][2 n_slots + n_tcols]
    double* sums;            // [n_eval][3]
    int n_slots, n_tcols, n_cta, want_l2;
};

__global__ void residual_final_kernel(ResidFinalParams p) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= p.n_slots) return;
    const int nsum = 2 * p.n_slots + p.n_tcols;
    const int tci = p.slot_tci[s];
    double s1 = 0.0, s2 = 0.0, st = 0.0;
    for (int b = 0; b < p.n_cta; ++b) {
        const double* q = p.partial + (size_t)b * nsum;
        s1 += q[s];
        s2 += q[p.n_slots + s];
        st += q[2 * p.n_slots + tci];
    }
    const int ev = p.slot_ev[s];
    p.sums[(long long)ev * 3 + 0] = s1;
    p.sums[(long long)ev * 3 + 1] = st;
    p.sums[(long long)ev * 3 + 2] = p.want_l2 ? s2 : nan("");
}

// ------------------------------------------------------------------------------------------
// host side: window plan (supplementary.py:411-425), contexts, batches
// ------------------------------------------------------------------------------------------
struct AxisPlan {
    long long n_global = 0, offset = 0, n_local = 0;
    long long window = 0, step = 1;
    std::vector<long long> starts, bounds;     // global coordinates
    std::vector<int> seg_lo, seg_hi;           // per window: segment range [lo, hi)
    int nseg = 0;
    // local cells: segments intersecting [offset, offset + n_local)
    struct Cell { int seg; int lo, hi; };       // local index range
    std::vector<Cell> cells;

    void build(long long ng, long long off, long long nl) {
        n_global = ng; offset = off; n_local = nl;
        window = ng / 2;
        const long long horizons = window + 1;
        step = std::max<long long>(1, horizons / 30);
        starts.clear();
        for (long long s = 0; s < horizons; s += step) starts.push_back(s);
        std::set<long long> edges;
        edges.insert(0); edges.insert(ng);
        for (long long s : starts) { edges.insert(s); edges.insert(s + window); }
        bounds.assign(edges.begin(), edges.end());
        nseg = (int)bounds.size() - 1;
        std::map<long long, int> pos;
        for (int i = 0; i < (int)bounds.size(); ++i) pos[bounds[i]] = i;
        seg_lo.clear(); seg_hi.clear();
        for (long long s : starts) { seg_lo.push_back(pos[s]); seg_hi.push_back(pos[s + window]); }
        cells.clear();
        for (int s = 0; s < nseg; ++s) {
            const long long lo = std::max(bounds[s], off), hi = std::min(bounds[s + 1], off + nl);
            if (hi > lo) cells.push_back({s, (int)(lo - off), (int)(hi - off)});
        }
    }
};

// split cells longer than `maxlen` into sub-cells (same segment id)
static std::vector<AxisPlan::Cell> split_cells(const std::vector<AxisPlan::Cell>& in, int maxlen) {
    std::vector<AxisPlan::Cell> out;
    for (auto& c : in)
        for (int lo = c.lo; lo < c.hi; lo += maxlen) out.push_back({c.seg, lo, std::min(c.hi, lo + maxlen)});
    return out;
}

struct DevBuf {
    void* ptr = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (ptr) { cudaFree(ptr); ptr = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&ptr, want);
        if (e != cudaSuccess) {
            ptr = nullptr;
            return fail("cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
        }
        cap = want;
        return 0;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = 0; }
};

struct epde_ctx {
    int device = 0;
    int ndim = 0;
    long long shape[EPDE_MAX_NDIM] = {0, 0, 0};
    long long N = 0;
    cudaStream_t stream = 0;
    std::map<int, double*> tokens;
    double* weights = nullptr;   // sample weights g (device), null = all ones
    AxisPlan axes[EPDE_MAX_NDIM];
    bool geom_ready = false;
    Geom geom;
    int n_cells = 0;             // nA * nB * npieces
    int tile_cols = 32;          // pass-A unit width limit = last-axis extent of the token tile / TMA box
    int box_a = 1, box_b = 1;    // leading extents of the token tile / TMA box (largest cell of slot A / slot B)
    int box_c = 32;              // its last-axis extent: tile_cols, or tile_cols + 2 when units may start on odd columns
    DevBuf d_geom;               // packed geometry tables
    DevBuf d_cells, d_seg, d_win, d_enc, d_out, d_partial, d_given, d_jobs, d_fd, d_total, d_marg;
    void* h_pinned = nullptr; size_t h_pinned_cap = 0;
    epde_batch* live = nullptr;
    long long launches = 0;
    size_t cell_budget_bytes = (size_t)16 << 30;   // scratch cap for the cell table of one launch group
    int sm_count = 148;
    int max_smem = 227 * 1024;   // opt-in shared memory per block of the device
    int smem_gram = 0, smem_resid = 0, smem_solve = 0;   // dynamic shared memory available to each kernel
    // optional per-kernel-class device timing (epde_ctx_set_timing)
    bool timing = false;
    struct Span { int cls; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> free_events;
    double class_ms[EPDE_N_KERNEL_CLASSES] = {0, 0, 0, 0};
    long long class_launches[EPDE_N_KERNEL_CLASSES] = {0, 0, 0, 0};
};

static cudaEvent_t take_event(epde_ctx* c) {
    cudaEvent_t e = nullptr;
    if (!c->free_events.empty()) { e = c->free_events.back(); c->free_events.pop_back(); }
    else cudaEventCreate(&e);
    return e;
}

// RAII bracket of one kernel launch: two event records on the context stream when timing is on
struct KernelSpan {
    epde_ctx* c; int cls; cudaEvent_t a = nullptr;
    KernelSpan(epde_ctx* c_, int cls_) : c(c_), cls(cls_) {
        c->launches++;
        if (c->timing) { a = take_event(c); cudaEventRecord(a, c->stream); }
    }
    ~KernelSpan() {
        if (a) { cudaEvent_t b = take_event(c); cudaEventRecord(b, c->stream); c->spans.push_back({cls, a, b}); }
    }
};

static int drain_spans(epde_ctx* c) {
    if (c->spans.empty()) return 0;
    CK(cudaStreamSynchronize(c->stream));
    for (auto& sp : c->spans) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, sp.a, sp.b));
        c->class_ms[sp.cls] += ms;
        c->class_launches[sp.cls]++;
        c->free_events.push_back(sp.a);
        c->free_events.push_back(sp.b);
    }
    c->spans.clear();
    return 0;
}

static int ensure_pinned(epde_ctx* c, size_t bytes) {
    if (bytes <= c->h_pinned_cap) return 0;
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    c->h_pinned = nullptr; c->h_pinned_cap = 0;
    CK(cudaMallocHost(&c->h_pinned, bytes + 4096));
    c->h_pinned_cap = bytes + 4096;
    return 0;
}

static int build_geometry(epde_ctx* c) {
    const int nd = c->ndim;
    // slot A / slot B = leading axes (dummy when absent), last axis = columns
    const int last = nd - 1;
    std::vector<AxisPlan::Cell> cellsA, cellsB;
    long long strideA = 0, strideB = 0;
    if (nd == 3) {
        cellsB = split_cells(c->axes[1].cells, TILE_ROWS);
        int widestB = 1;
        for (auto& cl : cellsB) widestB = std::max(widestB, cl.hi - cl.lo);
        cellsA = split_cells(c->axes[0].cells, std::max(1, TILE_ROWS / widestB));
        strideA = c->shape[1] * c->shape[2]; strideB = c->shape[2];
    } else if (nd == 2) {
        cellsA = {{0, 0, 1}};
        cellsB = split_cells(c->axes[0].cells, TILE_ROWS);
        strideA = 0; strideB = c->shape[1];
    } else {
        cellsA = {{0, 0, 1}}; cellsB = {{0, 0, 1}};
    }
    const int nA = (int)cellsA.size(), nB = (int)cellsB.size();
    int max_rows = 1;
    for (int a = 0; a < nA; ++a)
        for (int b = 0; b < nB; ++b)
            max_rows = std::max(max_rows, (cellsA[a].hi - cellsA[a].lo) * (cellsB[b].hi - cellsB[b].lo));
    // unit width: about 512 points per unit (the token tile of a unit is max_rows x tile_cols doubles per token)
    int tile_cols = 32;
    while (tile_cols > 8 && max_rows * tile_cols > 512) tile_cols /= 2;
    const AxisPlan& lastp = c->axes[last];
    const long long nl = c->shape[last];
    // pieces: sub-segments of the last axis no wider than a unit
    std::vector<int> piece_start, piece_seg;
    for (auto& cl : lastp.cells)
        for (int lo = cl.lo; lo < cl.hi; lo += tile_cols) { piece_start.push_back(lo); piece_seg.push_back(cl.seg); }
    piece_start.push_back((int)nl);
    const int npieces = (int)piece_seg.size();
    // chunks: runs of pieces with total width <= tile_cols
    struct Chunk { int p_lo, np, col_lo, ncols; };
    std::vector<Chunk> chunks;
    bool odd = false;
    for (int pI = 0; pI < npieces;) {
        int q = pI;
        while (q < npieces && piece_start[q + 1] - piece_start[pI] <= tile_cols) ++q;
        chunks.push_back({pI, q - pI, piece_start[pI], piece_start[q] - piece_start[pI]});
        if (piece_start[pI] & 1) odd = true;
        pI = q;
    }
    std::vector<UnitDesc> units;
    units.reserve((size_t)nA * nB * chunks.size());
    for (int a = 0; a < nA; ++a)
        for (int b = 0; b < nB; ++b)
            for (auto& ch : chunks) {
                UnitDesc u;
                memset(&u, 0, sizeof u);
                u.a_lo = cellsA[a].lo; u.a_hi = cellsA[a].hi;
                u.b_lo = cellsB[b].lo; u.b_hi = cellsB[b].hi;
                u.col_lo = ch.col_lo; u.ncols = ch.ncols;
                u.cell = a * nB + b;
                u.piece_lo = ch.p_lo; u.npieces = ch.np;
                units.push_back(u);
            }
    // segment rows: axes stacked; marginal descriptors (cells of one segment are contiguous)
    std::vector<int> axis_row_base(nd + 1, 0);
    for (int d = 0; d < nd; ++d) axis_row_base[d + 1] = axis_row_base[d] + c->axes[d].nseg;
    const int n_rows = axis_row_base[nd];
    std::vector<MargDesc> marg(n_rows);
    auto seg_range = [](const std::vector<AxisPlan::Cell>& cl, int s, int& lo, int& hi, long long& width) {
        lo = -1; hi = -1; width = 0;
        for (int q = 0; q < (int)cl.size(); ++q)
            if (cl[q].seg == s) { if (lo < 0) lo = q; hi = q + 1; width += cl[q].hi - cl[q].lo; }
        if (lo < 0) { lo = 0; hi = 0; }
    };
    for (int d = 0; d < nd; ++d) {
        for (int s = 0; s < c->axes[d].nseg; ++s) {
            MargDesc m;
            memset(&m, 0, sizeof m);
            m.row = axis_row_base[d] + s;
            m.a_lo = 0; m.a_hi = nA; m.b_lo = 0; m.b_hi = nB; m.p_lo = 0; m.p_hi = npieces;
            long long width = 0;
            m.kind = (d == last) ? 2 : (nd == 3 && d == 0) ? 0 : 1;
            if (d == last) {
                int lo = -1, hi = -1;
                for (int q = 0; q < npieces; ++q)
                    if (piece_seg[q] == s) { if (lo < 0) lo = q; hi = q + 1; width += piece_start[q + 1] - piece_start[q]; }
                if (lo < 0) { lo = 0; hi = 0; }
                m.p_lo = lo; m.p_hi = hi;
            } else if (nd == 3 && d == 0) {
                seg_range(cellsA, s, m.a_lo, m.a_hi, width);
            } else {
                seg_range(cellsB, s, m.b_lo, m.b_hi, width);
            }
            m.count = (double)(width * (c->N / c->shape[d]));
            marg[m.row] = m;
        }
    }
    // windows: segment-row ranges; last entry = every row of axis 0 (all points)
    std::vector<int> win_lo, win_hi;
    for (int d = 0; d < nd; ++d)
        for (size_t w = 0; w < c->axes[d].starts.size(); ++w) {
            win_lo.push_back(axis_row_base[d] + c->axes[d].seg_lo[w]);
            win_hi.push_back(axis_row_base[d] + c->axes[d].seg_hi[w]);
        }
    const int W = (int)win_lo.size();
    win_lo.push_back(axis_row_base[0]);
    win_hi.push_back(axis_row_base[1]);
    // pack and upload
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t o_units = 0;
    size_t o_piece = align(o_units + units.size() * sizeof(UnitDesc));
    size_t o_marg = align(o_piece + piece_start.size() * sizeof(int));
    size_t o_wlo = align(o_marg + marg.size() * sizeof(MargDesc));
    size_t o_whi = align(o_wlo + win_lo.size() * sizeof(int));
    size_t total = align(o_whi + win_hi.size() * sizeof(int));
    std::vector<char> host(total, 0);
    memcpy(host.data() + o_units, units.data(), units.size() * sizeof(UnitDesc));
    memcpy(host.data() + o_piece, piece_start.data(), piece_start.size() * sizeof(int));
    memcpy(host.data() + o_marg, marg.data(), marg.size() * sizeof(MargDesc));
    memcpy(host.data() + o_wlo, win_lo.data(), win_lo.size() * sizeof(int));
    memcpy(host.data() + o_whi, win_hi.data(), win_hi.size() * sizeof(int));
    if (c->d_geom.ensure(total)) return 1;
    CK(cudaMemcpy(c->d_geom.ptr, host.data(), total, cudaMemcpyHostToDevice));
    char* base = (char*)c->d_geom.ptr;
    Geom& g = c->geom;
    g.strideA = strideA; g.strideB = strideB;
    g.nA = nA; g.nB = nB; g.npieces = npieces;
    g.n_units = (int)units.size();
    g.n_cells = nA * nB * npieces;
    g.n_rows = n_rows;
    g.n_windows = W;
    g.ndim = nd;
    g.max_rows = max_rows;
    g.aligned16 = (nd == 1 || nl % 2 == 0) ? 1 : 0;
    g.N = c->N;
    g.units = (const UnitDesc*)(base + o_units);
    g.piece_start = (const int*)(base + o_piece);
    g.marg = (const MargDesc*)(base + o_marg);
    g.win_lo_row = (const int*)(base + o_wlo);
    g.win_hi_row = (const int*)(base + o_whi);
    c->n_cells = nA * nB * npieces;
    // the TMA box starts on an even column (16-byte aligned global address): a unit that starts on an odd column
    // fetches one column more on the left, so the box is two columns wider when such units exist
    c->tile_cols = tile_cols;
    c->box_c = odd ? tile_cols + 2 : tile_cols;
    c->box_a = 1; c->box_b = 1;
    for (auto& cl : cellsA) c->box_a = std::max(c->box_a, cl.hi - cl.lo);
    for (auto& cl : cellsB) c->box_b = std::max(c->box_b, cl.hi - cl.lo);
    c->geom_ready = true;
    return 0;
}

struct LaunchGroup {             // equations evaluated together over ONE dictionary of distinct term columns
    std::vector<int> eqs;
    std::vector<int> token_ids;  // slot -> token id
    std::vector<int> col_term;   // column -> global distinct-term id (the ones column is appended after these)
    int ncols = 0, cp = 0, n_ent = 0;
    int ntasks = 0, ksplit = 1, n_block_pairs = 0;
    int n_slots = 0, n_tcols = 0;   // pass-B jobs: evals (sum slots), distinct target columns
    int n_gfac = 0;
    long long seg_off = 0, win_off = 0, pe_off = 0;     // doubles / doubles / ints into the batch-wide tables
    // device pointers into ctx->d_enc
    ColDesc* d_cols = nullptr; FactorDesc* d_gfac = nullptr; TaskDesc* d_tasks = nullptr;
    int* d_flush = nullptr; int* d_slot_ev = nullptr; int* d_slot_tci = nullptr; int* d_tcols = nullptr;
};

struct epde_batch {
    epde_ctx* ctx = nullptr;
    int n_eq = 0, n_eval = 0;
    std::vector<EqDesc> eqs;
    std::vector<LaunchGroup> groups;
    std::vector<int> evals_gs8, evals_gs16, evals_gs32;
    // device pointers into ctx->d_enc
    EqDesc* d_eqs = nullptr;
    int* d_eval_eq = nullptr; int* d_eval_tgt = nullptr;
    int* d_lists = nullptr;                      // eval lists by GS
    int* d_pairent = nullptr;
    int off_gs8 = 0, off_gs16 = 0, off_gs32 = 0;
    // outputs in ctx->d_out
    double* d_coef = nullptr; double* d_intercept = nullptr; double* d_stab = nullptr;
    double* d_sw = nullptr; double* d_wterm = nullptr; double* d_sums = nullptr; double* d_b1 = nullptr;
    double* d_nact = nullptr; double* d_allzero = nullptr; double* d_iters = nullptr;
    size_t out_bytes = 0;      // solution block + residual sums: what epde_batch_fetch always copies
    size_t sol_bytes = 0;      // solution block (everything the solve writes except the window weights)
    size_t sw_bytes = 0;
    long long seg_doubles = 0, win_doubles = 0;
    int* d_mode = nullptr; double* d_given_w = nullptr; double* d_given_b = nullptr;
    bool solved = false;
    bool want_l2 = true;       // epde_batch_set_norms
    bool weighted = false;     // the context held non-trivial sample weights when the batch was created
};

// ---- pass-A tasks of a launch group: TASK_A x TASK_B rectangles of 8x8 column-block pairs covering the needed part
//      of the upper block triangle; ordered so that the four warp schedulers of an SM get equal DMMA counts
static void build_tasks(int nblk, const std::vector<char>& need /* [nblk * nblk], I <= J */, std::vector<TaskDesc>& tasks) {
    tasks.clear();
    for (int i0 = 0; i0 < nblk; i0 += TASK_A) {
        std::vector<int> js;
        for (int j = i0; j < nblk; ++j) {
            bool any = false;
            for (int i = i0; i < std::min(nblk, i0 + TASK_A); ++i)
                if (i <= j && need[i * nblk + j]) any = true;
            if (any) js.push_back(j);
        }
        for (size_t q = 0; q < js.size(); q += TASK_B) {
            TaskDesc t;
            memset(&t, 0, sizeof t);
            int pairs = 0;
            for (int a = 0; a < TASK_A; ++a) t.ablk[a] = (unsigned char)std::min(nblk - 1, i0 + a);
            for (int b = 0; b < TASK_B; ++b) t.bblk[b] = (unsigned char)js[std::min(js.size() - 1, q + b)];
            for (int a = 0; a < TASK_A; ++a)
                for (int b = 0; b < TASK_B; ++b) {
                    const int i = i0 + a, j = (q + b < js.size()) ? js[q + b] : -1;
                    if (i < nblk && j >= 0 && i <= j && need[i * nblk + j]) { t.mask |= (unsigned char)(1u << (a * TASK_B + b)); ++pairs; }
                }
            t.pad0 = (unsigned char)pairs;
            if (pairs) tasks.push_back(t);
        }
    }
    // longest-processing-time deal into the 4 scheduler bins (warp w runs on scheduler w % 4)
    std::stable_sort(tasks.begin(), tasks.end(), [](const TaskDesc& a, const TaskDesc& b) { return a.pad0 > b.pad0; });
    std::vector<TaskDesc> bins[4];
    int load[4] = {0, 0, 0, 0};
    for (auto& t : tasks) {
        int k = 0;
        for (int q = 1; q < 4; ++q) if (load[q] < load[k]) k = q;
        bins[k].push_back(t); load[k] += t.pad0;
    }
    // warp w sits on scheduler w % 4 and takes bins[w % 4][w / 4]: bins with more tasks first, so that the interleaved
    // order has no holes
    std::stable_sort(std::begin(bins), std::end(bins), [](const std::vector<TaskDesc>& a, const std::vector<TaskDesc>& b) {
        return a.size() > b.size(); });
    std::vector<TaskDesc> out;
    for (size_t r = 0; out.size() < tasks.size(); ++r)
        for (int k = 0; k < 4; ++k) if (r < bins[k].size()) out.push_back(bins[k][r]);
    tasks.swap(out);
}

struct TermKey {
    std::vector<double> v;       // sorted (token, kind, power, param) tuples, flattened
    bool operator<(const TermKey& o) const { return v < o.v; }
};

// ------------------------------------------------------------------------------------------
// token producer on the device: finite-difference derivatives exactly as
// AdaptiveFiniteDeriv does them (preprocessing/deriv_calculators.py:46-102 = repeated numpy.gradient with the
// coordinate array of the axis, edge_order 1), on the FULL grid, then the interior crop that the boundary mask
// applies before a tensor enters the cache (cache.py:250-270, interface.py:133-155).
// numpy.gradient: constant spacing -> (f[i+1] - f[i-1]) / (2 dx); otherwise a f[i-1] + b f[i] + c f[i+1] with
// a = -dx2 / (dx1 (dx1 + dx2)), b = (dx2 - dx1) / (dx1 dx2), c = dx1 / (dx2 (dx1 + dx2)); first-order one-sided
// differences at the two ends.  Products and sums are rounded separately (no FMA) and in numpy's order, so the
// result is bit-identical to the reference's.
// ------------------------------------------------------------------------------------------
struct FdParams {
    const double* in;
    double* out;
    long long total, stride;     // elements, element stride of the axis
    int n;                       // points along the axis
    int uniform;
    double dx;                   // uniform spacing
    const double* a; const double* b; const double* c;   // [n] interior coefficients (non-uniform)
    double dx_first, dx_last;    // end spacings
};

// I = unsigned when the tensor has fewer than 2^31 elements: the per-element index arithmetic (two divisions) is what
// these elementwise kernels spend their instructions on, and 64-bit division costs ~5x the 32-bit one
template <typename I>
__global__ void __launch_bounds__(256) fd_gradient_kernel(FdParams p) {
    const I total = (I)p.total, stride = (I)p.stride, n = (I)p.n, step = (I)gridDim.x * (I)blockDim.x;
    for (I idx = (I)blockIdx.x * (I)blockDim.x + (I)threadIdx.x; idx < total; idx += step) {
        const int i = (int)((idx / stride) % n);
        double v;
        if (i == 0) {
            v = __ddiv_rn(__dsub_rn(p.in[idx + p.stride], p.in[idx]), p.dx_first);
        } else if (i == p.n - 1) {
            v = __ddiv_rn(__dsub_rn(p.in[idx], p.in[idx - p.stride]), p.dx_last);
        } else if (p.uniform) {
            v = __ddiv_rn(__dsub_rn(p.in[idx + p.stride], p.in[idx - p.stride]), __dmul_rn(2.0, p.dx));
        } else {
            const double t0 = __dmul_rn(p.a[i], p.in[idx - p.stride]);
            const double t1 = __dmul_rn(p.b[i], p.in[idx]);
            const double t2 = __dmul_rn(p.c[i], p.in[idx + p.stride]);
            v = __dadd_rn(__dadd_rn(t0, t1), t2);
        }
        p.out[idx] = v;
    }
}

struct CropParams {
    const double* in;
    double* out;
    long long n_inner;
    int ndim;
    long long inner[EPDE_MAX_NDIM], full_stride[EPDE_MAX_NDIM], off[EPDE_MAX_NDIM];
};

template <typename I>
__global__ void __launch_bounds__(256) crop_interior_kernel(CropParams p) {
    const I total = (I)p.n_inner, step = (I)gridDim.x * (I)blockDim.x;
    for (I idx = (I)blockIdx.x * (I)blockDim.x + (I)threadIdx.x; idx < total; idx += step) {
        I rem = idx, src = 0;
        for (int d = p.ndim - 1; d >= 0; --d) {
            const I in_d = (I)p.inner[d];
            const I q = rem % in_d;
            rem /= in_d;
            src += (q + (I)p.off[d]) * (I)p.full_stride[d];
        }
        p.out[idx] = p.in[src];
    }
}

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------

// Chebyshev-window derivative (preprocessing/cheb.py:14-85 as PolynomialDeriv applies it point by point,
// deriv_calculators.py:111-134): at an interior index of the axis the m-th derivative of the least-squares Chebyshev
// polynomial through the `window` neighbouring values, evaluated at the centre, is a LINEAR functional of those values;
// the host computes its weights per index with numpy's own Chebyshev.fit (epde_b200/producers.py), the kernel applies
// them.  Within `bound` of either end the reference switches to one-sided finite differences (preprocessing/fin_der.py).
struct PolyParams {
    const double* in;
    double* out;
    const double* w;         // [n][window] weights of the derivative at index i over in[i - window/2 .. i + window/2]
    const double* x;         // [n] coordinates of the axis
    long long total, stride; // elements; distance between neighbours along the axis
    int n, window, bound, order;
};

template <typename I>
__global__ void __launch_bounds__(256) poly_derivative_kernel(PolyParams p) {
    const I total = (I)p.total, stride = (I)p.stride, np_ = (I)p.n, step = (I)gridDim.x * (I)blockDim.x;
    for (I q = (I)blockIdx.x * (I)blockDim.x + (I)threadIdx.x; q < total; q += step) {
        const int i = (int)((q / stride) % np_);
        const double* c = p.in + q;
        double v;
        if (i >= p.bound && i <= p.n - p.bound) {
            const int h = (p.window - 1) / 2;
            const double* w = p.w + (size_t)i * p.window;
            v = 0.0;
            for (int j = 0; j < p.window; ++j) v = fma(w[j], c[(long long)(j - h) * p.stride], v);
        } else {
            const int sgn = (i < p.bound) ? 1 : -1;             // fin_der.py:13-17: I = idx +/- arange(6)
            const double dx = p.x[i + sgn] - p.x[i];
            const double f0 = c[0], f1 = c[(long long)sgn * p.stride], f2 = c[(long long)2 * sgn * p.stride],
                         f3 = c[(long long)3 * sgn * p.stride], f4 = c[(long long)4 * sgn * p.stride];
            if (p.order == 1) v = (f1 - f0) / dx;
            else if (p.order == 2) v = (2 * f0 - 5 * f1 + 4 * f2 - f3) / (dx * dx);
            else v = (-2.5 * f0 + 9 * f1 - 12 * f2 + 7 * f3 - 1.5 * f4) / (dx * dx * dx);
        }
        p.out[q] = v;
    }
}

// sustained FP64 rate measurement (epde_measure_fp64_peak): independent FMA chains / DMMA chains, no memory traffic
template <int CHAINS>
__global__ void __launch_bounds__(256) fp64_dfma_kernel(double* out, int iters, double a, double b) {
    double acc[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) acc[k] = threadIdx.x * 1e-3 + k;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) acc[k] = fma(acc[k], a, b);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS>
__global__ void __launch_bounds__(256) fp64_dmma_kernel(double* out, int iters, double a, double b) {
    double d0[CHAINS], d1[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) { d0[k] = threadIdx.x * 1e-3 + k; d1[k] = k; }
    const double av = a + threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) dmma884(d0[k], d1[k], av, b);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += d0[k] + d1[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
extern "C" {

const char* epde_last_error(void) { return g_err.c_str(); }
const char* epde_version(void) { return "epde_b200 0.3.0 sm_100a"; }

static int set_dyn_smem(const void* fn, int max_smem, int* dyn_out) {
    cudaFuncAttributes fa;
    cudaError_t fe = cudaFuncGetAttributes(&fa, fn);
    int dyn = max_smem;
    if (fe == cudaSuccess) {
        dyn = max_smem - (int)fa.sharedSizeBytes;
        fe = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
    }
    if (fe != cudaSuccess)
        return fail("epde_ctx_create: cannot configure shared memory (%s); is this an sm_100 device?", cudaGetErrorString(fe));
    if (dyn_out) *dyn_out = std::min(*dyn_out, dyn);
    return 0;
}

int epde_ctx_create(int device, int ndim, const int64_t* inner_shape, epde_ctx** out) {
    if (!out || !inner_shape) return fail("epde_ctx_create: null argument");
    if (ndim < 1 || ndim > EPDE_MAX_NDIM) return fail("epde_ctx_create: ndim %d not in [1, %d]", ndim, EPDE_MAX_NDIM);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail("epde_ctx_create: no CUDA device available (%s); this engine has no CPU fallback",
                    cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail("epde_ctx_create: device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    epde_ctx* c = new epde_ctx();
    c->device = device;
    c->ndim = ndim;
    c->sm_count = prop.multiProcessorCount;
    c->max_smem = (int)prop.sharedMemPerBlockOptin;
    {
        // dynamic shared memory limit of a kernel = opt-in maximum minus its static shared memory
        c->smem_gram = c->smem_resid = c->smem_solve = c->max_smem;
        const void* gram[] = {(const void*)gram_dmma_kernel<false>, (const void*)gram_dmma_kernel<true>};
        const void* resid[] = {(const void*)residual_cols_kernel<false>, (const void*)residual_cols_kernel<true>};
        const void* solve[] = {(const void*)lasso_solve_kernel<8>, (const void*)lasso_solve_kernel<16>, (const void*)lasso_solve_kernel<32>};
        for (const void* fn : gram) if (set_dyn_smem(fn, c->max_smem, &c->smem_gram)) { delete c; return 1; }
        for (const void* fn : resid) if (set_dyn_smem(fn, c->max_smem, &c->smem_resid)) { delete c; return 1; }
        for (const void* fn : solve) if (set_dyn_smem(fn, c->max_smem, &c->smem_solve)) { delete c; return 1; }
        int unused = c->max_smem;
        if (set_dyn_smem((const void*)residual_pack_kernel, c->max_smem, &unused)) { delete c; return 1; }
        if (set_dyn_smem((const void*)marginal_slab_kernel, c->max_smem, &unused)) { delete c; return 1; }
    }
    c->N = 1;
    for (int d = 0; d < ndim; ++d) {
        if (inner_shape[d] < 2) { delete c; return fail("epde_ctx_create: axis %d has %lld points (< 2)", d, (long long)inner_shape[d]); }
        c->shape[d] = inner_shape[d];
        c->N *= inner_shape[d];
    }
    for (int d = 0; d < ndim; ++d) c->axes[d].build(c->shape[d], 0, c->shape[d]);
    if (build_geometry(c)) { delete c; return 1; }
    *out = c;
    return 0;
}

int epde_ctx_destroy(epde_ctx* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (auto& kv : c->tokens) cudaFree(kv.second);
    if (c->weights) cudaFree(c->weights);
    c->d_geom.release(); c->d_cells.release(); c->d_seg.release(); c->d_win.release(); c->d_enc.release();
    c->d_out.release(); c->d_partial.release(); c->d_given.release(); c->d_jobs.release(); c->d_fd.release();
    c->d_total.release(); c->d_marg.release();
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    for (auto& sp : c->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto& ev : c->free_events) cudaEventDestroy(ev);
    if (c->live) { c->live->ctx = nullptr; }
    delete c;
    return 0;
}

int epde_ctx_set_stream(epde_ctx* c, void* s) {
    if (!c) return fail("null context");
    c->stream = (cudaStream_t)s;
    return 0;
}

int epde_ctx_set_shard(epde_ctx* c, int64_t off, int64_t glob) {
    if (!c) return fail("null context");
    if (c->ndim < 2) return fail("epde_ctx_set_shard: grid sharding needs ndim >= 2");
    if (off < 0 || off + c->shape[0] > glob) return fail("epde_ctx_set_shard: rows [%lld, %lld) outside global axis of %lld",
                                                        (long long)off, (long long)(off + c->shape[0]), (long long)glob);
    CK(cudaSetDevice(c->device));
    c->axes[0].build(glob, off, c->shape[0]);
    return build_geometry(c);
}

// token vectors are allocated with 16 spare doubles: the 16-byte bulk copies of the streaming passes may read one
// element past the last point of an odd-length chunk
static int token_alloc(epde_ctx* c, double** p, const char* who) {
    if (*p) return 0;
    cudaError_t e = cudaMalloc((void**)p, sizeof(double) * ((size_t)c->N + 16));
    if (e != cudaSuccess) { *p = nullptr; return fail("%s: out of device memory (%s)", who, cudaGetErrorString(e)); }
    e = cudaMemsetAsync(*p + c->N, 0, sizeof(double) * 16, c->stream);
    if (e != cudaSuccess) return fail("%s: %s", who, cudaGetErrorString(e));
    return 0;
}

int epde_cache_upload(epde_ctx* c, int id, const double* host, int64_t n) {
    if (!c || !host) return fail("epde_cache_upload: null argument");
    if (n != c->N) return fail("epde_cache_upload: token %d has %lld values, context expects %lld", id, (long long)n, c->N);
    if (id < 0 || id >= EPDE_MAX_TOKENS) return fail("epde_cache_upload: token id %d out of range", id);
    CK(cudaSetDevice(c->device));
    double*& p = c->tokens[id];
    if (token_alloc(c, &p, "epde_cache_upload")) { c->tokens.erase(id); return 1; }
    CK(cudaMemcpyAsync(p, host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int epde_cache_upload_device(epde_ctx* c, int id, const double* dev, int64_t n) {
    if (!c || !dev) return fail("epde_cache_upload_device: null argument");
    if (n != c->N) return fail("epde_cache_upload_device: token %d has %lld values, context expects %lld", id, (long long)n, c->N);
    if (id < 0 || id >= EPDE_MAX_TOKENS) return fail("epde_cache_upload_device: token id %d out of range", id);
    CK(cudaSetDevice(c->device));
    double*& p = c->tokens[id];
    if (token_alloc(c, &p, "epde_cache_upload_device")) { c->tokens.erase(id); return 1; }
    CK(cudaMemcpyAsync(p, dev, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}

int epde_cache_download(epde_ctx* c, int id, double* host, int64_t n) {
    if (!c || !host) return fail("epde_cache_download: null argument");
    auto it = c->tokens.find(id);
    if (it == c->tokens.end()) return fail("epde_cache_download: token %d not in cache", id);
    if (n != c->N) return fail("epde_cache_download: size mismatch");
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(host, it->second, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int epde_cache_contains(epde_ctx* c, int id, int* present) {
    if (!c || !present) return fail("null argument");
    *present = c->tokens.count(id) ? 1 : 0;
    return 0;
}

int epde_cache_delete(epde_ctx* c, int id) {
    if (!c) return fail("null context");
    auto it = c->tokens.find(id);
    if (it == c->tokens.end()) return fail("epde_cache_delete: token %d not in cache", id);
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(it->second);
    c->tokens.erase(it);
    return 0;
}

int epde_cache_bytes(epde_ctx* c, int64_t* bytes) {
    if (!c || !bytes) return fail("null argument");
    *bytes = (int64_t)c->tokens.size() * c->N * (int64_t)sizeof(double);
    return 0;
}

int epde_ctx_set_sample_weights(epde_ctx* c, const double* host, int64_t n) {
    if (!c || !host) return fail("null argument");
    if (n != c->N) return fail("epde_ctx_set_sample_weights: size mismatch");
    CK(cudaSetDevice(c->device));
    bool ones = true;
    for (int64_t i = 0; i < n && ones; ++i) ones = host[i] == 1.0;
    if (ones) {                                 // the default 0/1 mask restricted to the interior (interface.py:540-545)
        if (c->weights) { CK(cudaStreamSynchronize(c->stream)); cudaFree(c->weights); c->weights = nullptr; }
        return 0;
    }
    if (token_alloc(c, &c->weights, "epde_ctx_set_sample_weights")) return 1;
    CK(cudaMemcpyAsync(c->weights, host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int epde_ctx_num_windows(epde_ctx* c, int* w) {
    if (!c || !w) return fail("null argument");
    *w = c->geom.n_windows;
    return 0;
}

int epde_ctx_num_segments(epde_ctx* c, int* s) {
    if (!c || !s) return fail("null argument");
    *s = c->geom.n_rows;
    return 0;
}

int epde_ctx_launch_count(epde_ctx* c, int64_t* l) {
    if (!c || !l) return fail("null argument");
    *l = c->launches;
    return 0;
}

int epde_ctx_set_timing(epde_ctx* c, int enabled) {
    if (!c) return fail("null context");
    CK(cudaSetDevice(c->device));
    if (drain_spans(c)) return 1;
    c->timing = enabled != 0;
    for (int k = 0; k < EPDE_N_KERNEL_CLASSES; ++k) { c->class_ms[k] = 0.0; c->class_launches[k] = 0; }
    return 0;
}

int epde_ctx_kernel_time(epde_ctx* c, int cls, double* ms, int64_t* launches) {
    if (!c) return fail("null context");
    if (cls < 0 || cls >= EPDE_N_KERNEL_CLASSES) return fail("epde_ctx_kernel_time: unknown kernel class %d", cls);
    CK(cudaSetDevice(c->device));
    if (drain_spans(c)) return 1;
    if (ms) *ms = c->class_ms[cls];
    if (launches) *launches = c->class_launches[cls];
    return 0;
}

int epde_batch_create(epde_ctx* c, int n_eq, const int32_t* term_off, const int32_t* fac_off,
                      const int32_t* fac_token, const double* fac_power, const int32_t* fac_kind,
                      const double* fac_param, int n_eval, const int32_t* eval_eq,
                      const int32_t* eval_target, epde_batch** out) {
    if (!c || !term_off || !fac_off || !fac_token || !fac_power || !out) return fail("epde_batch_create: null argument");
    if (n_eq <= 0) return fail("epde_batch_create: empty batch");
    if (n_eval < 0 || (n_eval > 0 && (!eval_eq || !eval_target))) return fail("epde_batch_create: bad eval list");
    if (c->live) return fail("epde_batch_create: another batch is live on this context");
    CK(cudaSetDevice(c->device));
    epde_batch* b = new epde_batch();
    b->ctx = c; b->n_eq = n_eq; b->n_eval = n_eval;
    b->weighted = c->weights != nullptr;
    const int n_terms = term_off[n_eq];
    const int n_fac = fac_off[n_terms];
    b->eqs.resize(n_eq);
    const int W = c->geom.n_windows;
    const int max_cols = CMAX - 1;             // term columns of a group (one more for the ones column)
    const int max_tok = b->weighted ? MAXU - 1 : MAXU;
    for (int f = 0; f < n_fac; ++f) {
        if (!c->tokens.count(fac_token[f])) { delete b; return fail("factor %d references token %d which is not in the cache", f, fac_token[f]); }
        const int kind = fac_kind ? fac_kind[f] : 0;
        if (kind < 0 || kind > 3) { delete b; return fail("factor %d has unknown kind %d", f, kind); }
    }
    // ---- dictionary of distinct terms: a term = the multiset of its (token, kind, power, param) factors
    std::map<TermKey, int> term_ids;
    std::vector<int> term_first;               // distinct term -> one (global) term index holding it
    std::vector<int> term_of(n_terms);         // (global) term index -> distinct term
    std::vector<std::vector<int>> term_tok;    // distinct term -> tokens
    for (int e = 0; e < n_eq; ++e) {
        const int T = term_off[e + 1] - term_off[e];
        if (T < 1 || T > TMAX) { delete b; return fail("equation %d has %d terms; supported: 1..%d", e, T, TMAX); }
        for (int t = term_off[e]; t < term_off[e + 1]; ++t) {
            const int nf = fac_off[t + 1] - fac_off[t];
            if (nf < 1 || nf > FMAXF) { delete b; return fail("equation %d term %d has %d factors; supported: 1..%d", e, t - term_off[e], nf, FMAXF); }
            std::vector<std::array<double, 4>> fs;
            for (int f = fac_off[t]; f < fac_off[t + 1]; ++f)
                fs.push_back({(double)fac_token[f], (double)(fac_kind ? fac_kind[f] : 0), fac_power[f], fac_param ? fac_param[f] : 0.0});
            std::sort(fs.begin(), fs.end());
            TermKey key;
            for (auto& q : fs) key.v.insert(key.v.end(), q.begin(), q.end());
            auto it = term_ids.find(key);
            if (it == term_ids.end()) {
                it = term_ids.emplace(key, (int)term_first.size()).first;
                term_first.push_back(t);
                std::set<int> tk;
                for (int f = fac_off[t]; f < fac_off[t + 1]; ++f) tk.insert(fac_token[f]);
                term_tok.push_back(std::vector<int>(tk.begin(), tk.end()));
            }
            term_of[t] = it->second;
        }
    }
    // ---- launch groups: equations (most terms first) go to the open group they share the most columns with,
    //      as long as its dictionary stays within CMAX columns and MAXU tokens
    std::vector<int> order(n_eq);
    for (int e = 0; e < n_eq; ++e) order[e] = e;
    std::stable_sort(order.begin(), order.end(), [&](int a, int bb) {
        return term_off[a + 1] - term_off[a] > term_off[bb + 1] - term_off[bb]; });
    std::vector<std::set<int>> g_terms, g_tok;
    for (int e : order) {
        std::set<int> mine, mtok;
        for (int t = term_off[e]; t < term_off[e + 1]; ++t) {
            mine.insert(term_of[t]);
            mtok.insert(term_tok[term_of[t]].begin(), term_tok[term_of[t]].end());
        }
        if ((int)mtok.size() > max_tok) { delete b; return fail("equation %d references %d distinct tokens; supported: %d", e, (int)mtok.size(), max_tok); }
        int gi = -1, best_new = 1 << 30;
        for (size_t k = 0; k < b->groups.size(); ++k) {
            int fresh = 0;
            for (int t : mine) fresh += g_terms[k].count(t) ? 0 : 1;
            std::set<int> uni = g_tok[k];
            uni.insert(mtok.begin(), mtok.end());
            if ((int)g_terms[k].size() + fresh <= max_cols && (int)uni.size() <= max_tok && fresh < best_new) { gi = (int)k; best_new = fresh; }
        }
        if (gi < 0) { b->groups.push_back(LaunchGroup()); g_terms.push_back({}); g_tok.push_back({}); gi = (int)b->groups.size() - 1; }
        LaunchGroup& g = b->groups[gi];
        for (int t = term_off[e]; t < term_off[e + 1]; ++t)
            if (g_terms[gi].insert(term_of[t]).second) g.col_term.push_back(term_of[t]);
        g_tok[gi].insert(mtok.begin(), mtok.end());
        g.eqs.push_back(e);
        b->eqs[e].group = gi;
    }
    // ---- evals
    std::vector<std::vector<int>> eq_evals(n_eq);
    {
        std::set<std::pair<int, int>> seen;
        for (int v = 0; v < n_eval; ++v) {
            const int e = eval_eq[v], t = eval_target[v];
            if (e < 0 || e >= n_eq) { delete b; return fail("eval %d references equation %d", v, e); }
            const int T = term_off[e + 1] - term_off[e];
            if (t < 0 || t >= T) { delete b; return fail("eval %d: target %d outside equation %d (%d terms)", v, t, e, T); }
            if (!seen.insert({e, t}).second) { delete b; return fail("eval %d duplicates (equation %d, target %d)", v, e, t); }
            eq_evals[e].push_back(v);
            if (T <= 8) b->evals_gs8.push_back(v); else if (T <= 16) b->evals_gs16.push_back(v); else b->evals_gs32.push_back(v);
        }
    }
    // ---- per group: columns, pair entries, tasks, flush maps, jobs -> one packed host image
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    std::vector<char> img;
    auto put = [&](const void* src, size_t bytes) { const size_t o = align(img.size()); img.resize(o + std::max<size_t>(bytes, 1)); if (bytes) memcpy(img.data() + o, src, bytes); return o; };
    struct GOff { size_t cols, gfac, tasks, flush, slot_ev, slot_tci, tcols; };
    std::vector<GOff> goff(b->groups.size());
    std::vector<int> pairent_all;
    long long seg_off = 0, win_off = 0;
    const size_t n_cells = (size_t)c->n_cells;
    for (size_t gi = 0; gi < b->groups.size(); ++gi) {
        LaunchGroup& g = b->groups[gi];
        g.token_ids.assign(g_tok[gi].begin(), g_tok[gi].end());
        std::map<int, int> slot;
        for (size_t u = 0; u < g.token_ids.size(); ++u) slot[g.token_ids[u]] = (int)u;
        // Column order: pass A computes whole 8 x 8 blocks of column pairs, and a block pair is needed as soon as one
        // equation has a column in each block, so columns that occur together should share a block.  Greedy fill: a
        // block is seeded with the unplaced column that co-occurs most with the other unplaced ones and takes, 7 times,
        // the unplaced column that shares the most equations with it (~10 % fewer block pairs than first-appearance
        // order on the bench populations; a 20 000-step local search finds 25 % but costs tens of ms on the host).
        if (g.col_term.size() > 8) {
            const int n = (int)g.col_term.size();
            std::map<int, int> idx;
            for (int k = 0; k < n; ++k) idx[g.col_term[k]] = k;
            std::vector<int> co((size_t)n * n, 0);
            for (int e : g.eqs) {
                std::vector<int> cs;
                for (int t = term_off[e]; t < term_off[e + 1]; ++t) cs.push_back(idx[term_of[t]]);
                for (int a : cs) for (int bb : cs) if (a != bb) co[(size_t)a * n + bb]++;
            }
            std::vector<char> placed(n, 0);
            std::vector<int> order_c;
            while ((int)order_c.size() < n) {
                int seed = -1; long best = -1;
                for (int cidx = 0; cidx < n; ++cidx) {
                    if (placed[cidx]) continue;
                    long sc = 0;
                    for (int u = 0; u < n; ++u) if (!placed[u]) sc += co[(size_t)cidx * n + u];
                    if (sc > best) { best = sc; seed = cidx; }
                }
                std::vector<int> blk{seed};
                placed[seed] = 1;
                while (blk.size() < 8 && (int)(order_c.size() + blk.size()) < n) {
                    int nxt = -1; long bsc = -1;
                    for (int cidx = 0; cidx < n; ++cidx) {
                        if (placed[cidx]) continue;
                        long sc = 0;
                        for (int q : blk) sc += co[(size_t)cidx * n + q];
                        if (sc > bsc) { bsc = sc; nxt = cidx; }
                    }
                    blk.push_back(nxt);
                    placed[nxt] = 1;
                }
                order_c.insert(order_c.end(), blk.begin(), blk.end());
            }
            // keep the clustered order only if it needs fewer block pairs than the order of first appearance (which is
            // already the best one when equations share nothing: each equation's columns sit next to each other)
            auto block_pairs_of = [&](const std::vector<int>& pos /* column index -> position */) {
                const int nb = (n + 1 + 7) / 8, ones_blk = n / 8;
                std::vector<char> need((size_t)nb * nb, 0);
                for (int e : g.eqs) {
                    std::vector<int> bl{ones_blk};
                    for (int t = term_off[e]; t < term_off[e + 1]; ++t) bl.push_back(pos[idx[term_of[t]]] / 8);
                    for (int a : bl) for (int bb : bl) need[(size_t)std::min(a, bb) * nb + std::max(a, bb)] = 1;
                }
                int cnt = 0;
                for (char v : need) cnt += v;
                return cnt;
            };
            std::vector<int> pos_first(n), pos_greedy(n);
            for (int k = 0; k < n; ++k) { pos_first[k] = k; pos_greedy[order_c[k]] = k; }
            if (block_pairs_of(pos_greedy) < block_pairs_of(pos_first)) {
                std::vector<int> re(n);
                for (int k = 0; k < n; ++k) re[k] = g.col_term[order_c[k]];
                g.col_term.swap(re);
            }
        }
        std::map<int, int> col_of;
        for (size_t k = 0; k < g.col_term.size(); ++k) col_of[g.col_term[k]] = (int)k;
        const int ones = (int)g.col_term.size();
        g.ncols = ones + 1;
        g.cp = (g.ncols + 7) & ~7;
        const int nblk = g.cp / 8;
        // columns
        const int nbuild = g.cp + (b->weighted ? 1 : 0);
        std::vector<ColDesc> cols(nbuild);
        std::vector<FactorDesc> gfac;
        for (int k = 0; k < nbuild; ++k) { memset(&cols[k], 0, sizeof(ColDesc)); cols[k].kind = 3; for (int q = 0; q < MAXSF; ++q) cols[k].off[q] = -1; }
        for (int k = 0; k < ones; ++k) {
            const int t = term_first[g.col_term[k]];
            ColDesc& cd = cols[k];
            bool fast = true;
            int nsf = 0;
            for (int f = fac_off[t]; f < fac_off[t + 1]; ++f) {
                const int kind = fac_kind ? fac_kind[f] : 0;
                const double pw = fac_power[f];
                const int ip = (pw == 1.0) ? 1 : (pw == 2.0) ? 2 : (pw == 3.0) ? 3 : 0;
                if (kind != 0 || ip == 0) { fast = false; continue; }
                for (int q = 0; q < ip; ++q) { if (nsf < MAXSF) cd.off[nsf] = slot[fac_token[f]]; ++nsf; }
            }
            if (nsf > MAXSF) fast = false;
            // plain integer powers that do not fit the expanded form: (slot, power) pairs, all in the descriptor
            bool powprod = !fast && (fac_off[t + 1] - fac_off[t]) <= 4;
            for (int f = fac_off[t]; powprod && f < fac_off[t + 1]; ++f) {
                const double pw = fac_power[f];
                if ((fac_kind ? fac_kind[f] : 0) != 0 || !(pw >= 1.0 && pw <= 255.0 && pw == (double)(int)pw)) powprod = false;
            }
            if (fast) { cd.kind = 0; cd.nf = (short)nsf; }
            else if (powprod) {
                cd.kind = 4; cd.nf = (short)(fac_off[t + 1] - fac_off[t]); cd.gfac_off = 0;
                for (int q = 0; q < MAXSF; ++q) cd.off[q] = -1;
                for (int f = fac_off[t], q = 0; f < fac_off[t + 1]; ++f, ++q) {
                    cd.off[q] = slot[fac_token[f]];
                    cd.gfac_off = (int)((unsigned)cd.gfac_off | ((unsigned)(int)fac_power[f] << (8 * q)));
                }
            }
            else {
                cd.kind = 1; cd.nf = (short)(fac_off[t + 1] - fac_off[t]); cd.gfac_off = (int)gfac.size();
                for (int f = fac_off[t]; f < fac_off[t + 1]; ++f) {
                    FactorDesc d;
                    memset(&d, 0, sizeof d);
                    d.slot = slot[fac_token[f]];
                    d.power = fac_power[f];
                    d.kind = fac_kind ? fac_kind[f] : 0;
                    d.param = fac_param ? fac_param[f] : 0.0;
                    d.ipow = (d.power >= 1.0 && d.power <= 64.0 && d.power == (double)(int)d.power) ? (int)d.power : 0;
                    gfac.push_back(d);
                }
            }
        }
        cols[ones].kind = 2;
        if (b->weighted) { cols[g.cp].kind = 0; cols[g.cp].nf = 1; cols[g.cp].off[0] = (int)g.token_ids.size(); }
        g.n_gfac = (int)gfac.size();
        // needed pairs -> entries
        std::vector<int> pairent((size_t)g.cp * g.cp, -1);
        for (int e : g.eqs) {
            EqDesc& q = b->eqs[e];
            memset(q.col, 0, sizeof q.col);
            q.T = term_off[e + 1] - term_off[e];
            for (int t = 0; t < q.T; ++t) q.col[t] = (short)col_of[term_of[term_off[e] + t]];
            q.col[q.T] = (short)ones;
            for (int i = 0; i <= q.T; ++i)
                for (int j = 0; j <= q.T; ++j) pairent[(size_t)q.col[i] * g.cp + q.col[j]] = 0;
        }
        int n_ent = 0;
        for (int i = 0; i < g.cp; ++i)
            for (int j = i; j < g.cp; ++j)
                if (pairent[(size_t)i * g.cp + j] == 0) { pairent[(size_t)i * g.cp + j] = pairent[(size_t)j * g.cp + i] = n_ent++; }
        g.n_ent = n_ent;
        if (n_cells * (size_t)n_ent * sizeof(double) > c->cell_budget_bytes) {
            delete b;
            return fail("launch group %zu needs a cell table of %zu bytes (> %zu): grid too fine for this many distinct terms",
                        gi, n_cells * (size_t)n_ent * sizeof(double), c->cell_budget_bytes);
        }
        // block pairs, tasks, flush maps
        std::vector<char> need((size_t)nblk * nblk, 0);
        for (int i = 0; i < g.cp; ++i)
            for (int j = i; j < g.cp; ++j)
                if (pairent[(size_t)i * g.cp + j] >= 0) need[(size_t)(i / 8) * nblk + j / 8] = 1;
        std::vector<TaskDesc> tasks;
        build_tasks(nblk, need, tasks);
        g.ntasks = (int)tasks.size();
        g.n_block_pairs = 0;
        for (auto& t : tasks) g.n_block_pairs += t.pad0;
        if (g.ntasks > GRAM_MAX_WARPS) { delete b; return fail("internal: %d pass-A tasks exceed %d warps", g.ntasks, GRAM_MAX_WARPS); }
        g.ksplit = (g.ntasks >= 8) ? 1 : std::max(1, std::min(4, GRAM_MAX_WARPS / g.ntasks));
        std::vector<int> flush((size_t)g.ntasks * TASK_PAIRS * 64, -1);
        for (int t = 0; t < g.ntasks; ++t)
            for (int m = 0; m < TASK_PAIRS; ++m) {
                if (!((tasks[t].mask >> m) & 1)) continue;
                const int I = tasks[t].ablk[m / TASK_B], J = tasks[t].bblk[m % TASK_B];
                for (int l = 0; l < 32; ++l)
                    for (int h = 0; h < 2; ++h) {
                        const int ci = I * 8 + l / 4, cj = J * 8 + 2 * (l % 4) + h;
                        if (ci <= cj) flush[((size_t)(t * TASK_PAIRS + m) * 32 + l) * 2 + h] = pairent[(size_t)ci * g.cp + cj];
                    }
            }
        // pass-B jobs.  Sum slots = the evals of the group sorted by target column (consecutive evals then share the
        // target registers); distinct target columns for sum|t|
        std::vector<std::pair<int, int>> byt;      // (target column, eval)
        for (int e : g.eqs)
            for (int v : eq_evals[e]) byt.push_back({(int)b->eqs[e].col[eval_target[v]], v});
        std::stable_sort(byt.begin(), byt.end());
        std::vector<int> slot_ev, slot_tci, tcols;
        for (auto& q : byt) {
            if (tcols.empty() || tcols.back() != q.first) tcols.push_back(q.first);
            slot_ev.push_back(q.second);
            slot_tci.push_back((int)tcols.size() - 1);
        }
        g.n_slots = (int)slot_ev.size();
        g.n_tcols = (int)tcols.size();
        // tables
        g.seg_off = seg_off; seg_off += (long long)c->geom.n_rows * n_ent + (b->weighted ? n_ent : 0);
        g.win_off = win_off; win_off += (long long)(W + 1) * n_ent;
        g.pe_off = (long long)pairent_all.size();
        pairent_all.insert(pairent_all.end(), pairent.begin(), pairent.end());
        for (int e : g.eqs) {
            EqDesc& q = b->eqs[e];
            q.ones_col = ones; q.n_ent = n_ent; q.w_off = g.win_off; q.pe_off = g.pe_off; q.cp = g.cp;
        }
        goff[gi].cols = put(cols.data(), cols.size() * sizeof(ColDesc));
        goff[gi].gfac = put(gfac.data(), gfac.size() * sizeof(FactorDesc));
        goff[gi].tasks = put(tasks.data(), tasks.size() * sizeof(TaskDesc));
        goff[gi].flush = put(flush.data(), flush.size() * sizeof(int));
        goff[gi].slot_ev = put(slot_ev.data(), slot_ev.size() * sizeof(int));
        goff[gi].slot_tci = put(slot_tci.data(), slot_tci.size() * sizeof(int));
        goff[gi].tcols = put(tcols.data(), tcols.size() * sizeof(int));
    }
    b->seg_doubles = seg_off;
    b->win_doubles = win_off;
    std::vector<int> lists;
    b->off_gs8 = (int)lists.size();  lists.insert(lists.end(), b->evals_gs8.begin(), b->evals_gs8.end());
    b->off_gs16 = (int)lists.size(); lists.insert(lists.end(), b->evals_gs16.begin(), b->evals_gs16.end());
    b->off_gs32 = (int)lists.size(); lists.insert(lists.end(), b->evals_gs32.begin(), b->evals_gs32.end());
    const size_t o_eqs = put(b->eqs.data(), sizeof(EqDesc) * n_eq);
    const size_t o_eveq = put(eval_eq, sizeof(int) * n_eval);
    const size_t o_evtg = put(eval_target, sizeof(int) * n_eval);
    const size_t o_list = put(lists.data(), sizeof(int) * lists.size());
    const size_t o_pe = put(pairent_all.data(), sizeof(int) * pairent_all.size());
    const size_t total = align(img.size());
    // the pinned buffer may still feed an earlier asynchronous copy
    cudaStreamSynchronize(c->stream);
    if (ensure_pinned(c, total) || c->d_enc.ensure(total)) { delete b; return 1; }
    memcpy(c->h_pinned, img.data(), img.size());
    cudaError_t ce = cudaMemcpyAsync(c->d_enc.ptr, c->h_pinned, total, cudaMemcpyHostToDevice, c->stream);
    if (ce != cudaSuccess) { delete b; return fail("encoding upload failed: %s", cudaGetErrorString(ce)); }
    char* d = (char*)c->d_enc.ptr;
    b->d_eqs = (EqDesc*)(d + o_eqs);
    b->d_eval_eq = (int*)(d + o_eveq);
    b->d_eval_tgt = (int*)(d + o_evtg);
    b->d_lists = (int*)(d + o_list);
    b->d_pairent = (int*)(d + o_pe);
    for (size_t gi = 0; gi < b->groups.size(); ++gi) {
        LaunchGroup& g = b->groups[gi];
        g.d_cols = (ColDesc*)(d + goff[gi].cols); g.d_gfac = (FactorDesc*)(d + goff[gi].gfac);
        g.d_tasks = (TaskDesc*)(d + goff[gi].tasks); g.d_flush = (int*)(d + goff[gi].flush);
        g.d_slot_ev = (int*)(d + goff[gi].slot_ev); g.d_slot_tci = (int*)(d + goff[gi].slot_tci); g.d_tcols = (int*)(d + goff[gi].tcols);
    }
    // outputs
    const size_t ne = (size_t)std::max(1, n_eval);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align(o + bytes); return r; };
    const size_t oc = take(ne * TMAX * 8), oi = take(ne * 8), ost = take(ne * 8), owt = take(ne * TMAX * 8), ob1 = take(ne * 8),
                 ona = take(ne * 8), oaz = take(ne * 8), oit = take(ne * 8);
    b->sol_bytes = o;
    const size_t osu = take(ne * 3 * 8);
    b->out_bytes = o;
    const size_t osw = take(ne * W * KS * 8);
    b->sw_bytes = o - osw;
    if (c->d_out.ensure(o)) { delete b; return 1; }
    char* od = (char*)c->d_out.ptr;
    b->d_coef = (double*)(od + oc); b->d_intercept = (double*)(od + oi); b->d_stab = (double*)(od + ost);
    b->d_sw = (double*)(od + osw); b->d_wterm = (double*)(od + owt); b->d_sums = (double*)(od + osu); b->d_b1 = (double*)(od + ob1);
    b->d_nact = (double*)(od + ona); b->d_allzero = (double*)(od + oaz); b->d_iters = (double*)(od + oit);
    c->live = b;
    *out = b;
    return 0;
}

int epde_batch_destroy(epde_batch* b) {
    if (!b) return 0;
    if (b->ctx) {
        cudaSetDevice(b->ctx->device);
        cudaStreamSynchronize(b->ctx->stream);
        if (b->ctx->live == b) b->ctx->live = nullptr;
    }
    delete b;
    return 0;
}

int epde_batch_set_given(epde_batch* b, const int32_t* mode, const double* wterm, const double* intercept) {
    if (!b || !b->ctx || !mode || !wterm || !intercept) return fail("epde_batch_set_given: null argument");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    const size_t ne = (size_t)std::max(1, b->n_eval);
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_w = align(ne * sizeof(int)), o_b = align(o_w + ne * TMAX * 8), total = align(o_b + ne * 8);
    for (int v = 0; v < b->n_eval; ++v)
        if (mode[v] < 0 || mode[v] > 2) return fail("epde_batch_set_given: eval %d has mode %d", v, mode[v]);
    if (c->d_given.ensure(total)) return 1;
    char* d = (char*)c->d_given.ptr;
    // pageable sources: the runtime stages them before returning, so the caller may reuse them
    CK(cudaMemcpyAsync(d, mode, sizeof(int) * b->n_eval, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d + o_w, wterm, sizeof(double) * b->n_eval * TMAX, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d + o_b, intercept, sizeof(double) * b->n_eval, cudaMemcpyHostToDevice, c->stream));
    b->d_mode = (int*)d; b->d_given_w = (double*)(d + o_w); b->d_given_b = (double*)(d + o_b);
    return 0;
}

int epde_batch_set_norms(epde_batch* b, int want_l2) {
    if (!b || !b->ctx) return fail("epde_batch_set_norms: null batch");
    b->want_l2 = want_l2 != 0;
    return 0;
}

int epde_batch_info(epde_batch* b, int64_t* info) {
    if (!b || !info) return fail("null argument");
    int64_t cols = 0, ent = 0, bp = 0, tasks = 0;
    for (auto& g : b->groups) {
        cols += g.ncols; ent += g.n_ent; bp += g.n_block_pairs; tasks += g.ntasks;
    }
    info[0] = (int64_t)b->groups.size(); info[1] = cols; info[2] = ent; info[3] = bp; info[4] = tasks; info[5] = b->weighted ? 1 : 0;
    return 0;
}

int epde_batch_gram_size(epde_batch* b, int64_t* n) {
    if (!b || !n) return fail("null argument");
    *n = b->seg_doubles;
    return 0;
}

int epde_batch_residual_size(epde_batch* b, int64_t* n) {
    if (!b || !n) return fail("null argument");
    *n = (int64_t)std::max(1, b->n_eval) * 3;
    return 0;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
static PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess) fn = (PFN_cuTensorMapEncodeTiled_v12000)f;
    }
    return fn;
}

// tensor map of one token vector seen as the context's N-d grid, box = one pass-A token tile
static int make_token_map(const epde_ctx* c, const double* base, CUtensorMap* out) {
    auto enc = tensor_map_encoder();
    if (!enc) return fail("cuTensorMapEncodeTiled is not available from this driver");
    const int nd = c->ndim;
    cuuint64_t dims[3]; cuuint64_t strides[2]; cuuint32_t box[3]; cuuint32_t estr[3] = {1, 1, 1};
    if (nd == 3) {
        dims[0] = c->shape[2]; dims[1] = c->shape[1]; dims[2] = c->shape[0];
        strides[0] = c->shape[2] * 8; strides[1] = c->shape[2] * c->shape[1] * 8;
        box[0] = c->box_c; box[1] = c->box_b; box[2] = c->box_a;
    } else if (nd == 2) {
        dims[0] = c->shape[1]; dims[1] = c->shape[0];
        strides[0] = c->shape[1] * 8;
        box[0] = c->box_c; box[1] = c->box_b;
    } else {
        dims[0] = c->shape[0];
        box[0] = c->box_c;
    }
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, nd, (void*)base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return 0;
}

// one pass-A launch of a launch group: cell table of the (weighted or unweighted) column Gram
static int launch_gram(epde_ctx* c, epde_batch* b, const LaunchGroup& g, double* cells, bool weighted) {
    (void)b;
    static GramParams p;                        // 2.2 KB with the tensor maps: keep it off the stack
    memset(&p, 0, sizeof p);
    p.g = c->geom; p.cols = g.d_cols; p.gfac = g.d_gfac;
    p.cp = g.cp;
    p.U = (int)g.token_ids.size();
    for (int u = 0; u < p.U; ++u) p.tok[u] = c->tokens[g.token_ids[u]];
    p.nbuild = g.cp; p.wcol = g.cp;
    if (weighted) { p.tok[p.U++] = c->weights; p.nbuild = g.cp + 1; }
    p.tasks = g.d_tasks; p.ntasks = g.ntasks; p.ksplit = g.ksplit; p.flushmap = g.d_flush;
    p.cells = cells; p.n_ent = g.n_ent;
    p.box_a = c->box_a; p.box_b = c->box_b; p.box_c = c->box_c;
    p.use_tma = (c->geom.aligned16 && tensor_map_encoder() && c->box_a <= 256 && c->box_b <= 256) ? 1 : 0;
    if (p.use_tma)
        for (int u = 0; u < p.U; ++u)
            if (make_token_map(c, p.tok[u], &p.tmap[u])) return 1;
    const int ncons = g.ntasks * g.ksplit;
    const int nwarps = ncons + std::min(GRAM_PROD_WARPS, GRAM_MAX_THREADS / 32 - ncons);
    const size_t tstride = ((size_t)p.box_a * p.box_b * p.box_c + 15) & ~(size_t)15;
    const size_t tile_bytes = sizeof(double) * (size_t)p.U * tstride;
    const size_t rest = sizeof(double) * 2 * (size_t)p.nbuild * PP_A + sizeof(ColDesc) * (size_t)p.nbuild +
                        (g.ksplit > 1 ? sizeof(double) * (size_t)g.ntasks * (g.ksplit - 1) * TASK_PAIRS * 2 * 32 : 0);
    p.nstages = (2 * tile_bytes + rest <= (size_t)c->smem_gram) ? 2 : 1;
    const size_t smem = p.nstages * tile_bytes + rest;
    if (smem > (size_t)c->smem_gram) return fail("pass A needs %zu bytes of shared memory (> %d)", smem, c->smem_gram);
    const unsigned grid = (unsigned)std::min<long long>(c->geom.n_units, (long long)c->sm_count);
    {
        KernelSpan ks(c, EPDE_KERNEL_GRAM);
        if (weighted) gram_dmma_kernel<true><<<grid, nwarps * 32, smem, c->stream>>>(p);
        else gram_dmma_kernel<false><<<grid, nwarps * 32, smem, c->stream>>>(p);
    }
    CK(cudaGetLastError());
    return 0;
}

int epde_batch_gram(epde_batch* b, double* dev_seg) {
    if (!b || !b->ctx) return fail("null batch");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    if (!dev_seg) {
        if (c->d_seg.ensure(sizeof(double) * (size_t)b->seg_doubles)) return 1;
        dev_seg = (double*)c->d_seg.ptr;
    }
    size_t need = 0;
    for (auto& g : b->groups) need = std::max(need, (size_t)c->n_cells * g.n_ent);
    if (c->d_cells.ensure(sizeof(double) * need)) return 1;
    double* cells = (double*)c->d_cells.ptr;
    for (auto& g : b->groups) {
        double* seg = dev_seg + g.seg_off;
        const unsigned ey = (unsigned)((g.n_ent + 127) / 128);
        if (b->weighted) {
            // unweighted total Gram first (LASSO sums, sparsity.py:54-55,73-76), kept behind the segment rows
            if (launch_gram(c, b, g, cells, false)) return 1;
            TotalParams tp;
            tp.cells = cells; tp.n_cells = c->n_cells; tp.n_ent = g.n_ent; tp.parts = std::min(c->n_cells, 256);
            if (c->d_total.ensure(sizeof(double) * (size_t)tp.parts * g.n_ent)) return 1;
            tp.part = (double*)c->d_total.ptr; tp.total = seg + (long long)c->geom.n_rows * g.n_ent;
            KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
            total_part_kernel<<<dim3(tp.parts, ey), 128, 0, c->stream>>>(tp);
            total_final_kernel<<<ey, 128, 0, c->stream>>>(tp);
            c->launches++;
            CK(cudaGetLastError());
        }
        if (launch_gram(c, b, g, cells, b->weighted)) return 1;
        const size_t slab_smem = sizeof(double) * (size_t)c->geom.npieces * 128;
        if (c->ndim == 3 && c->geom.nA >= 4 && slab_smem <= 100 * 1024) {
            MargSlabParams sp;
            sp.g = c->geom; sp.cells = cells; sp.seg = seg; sp.n_ent = g.n_ent;
            if (c->d_marg.ensure(sizeof(double) * (size_t)c->geom.nA * (MARG_NBC + c->geom.nB + MARG_NBC * c->geom.npieces) * g.n_ent))
                return 1;
            sp.part = (double*)c->d_marg.ptr;
            KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
            marginal_slab_kernel<<<dim3(c->geom.nA, ey, MARG_NBC), 128, slab_smem, c->stream>>>(sp);
            marginal_rows_kernel<<<dim3(c->geom.n_rows, ey), 128, 0, c->stream>>>(sp);
            c->launches++;
        } else {
            MargParams mp;
            mp.g = c->geom; mp.cells = cells; mp.seg = seg; mp.n_ent = g.n_ent;
            if (c->d_marg.ensure(sizeof(double) * (size_t)MARG_SPLIT * c->geom.n_rows * g.n_ent)) return 1;
            mp.part = (double*)c->d_marg.ptr;
            KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
            marginal_kernel<<<dim3(c->geom.n_rows, ey, MARG_SPLIT), 128, 0, c->stream>>>(mp);
            marginal_final_kernel<<<dim3(c->geom.n_rows, ey), 128, 0, c->stream>>>(mp);
            c->launches++;
        }
        CK(cudaGetLastError());
    }
    return 0;
}

static int solve_impl(epde_batch* b, const double* dev_seg, int part, int nparts) {
    if (!b || !b->ctx) return fail("null batch");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    const double* seg = dev_seg ? dev_seg : (const double*)c->d_seg.ptr;
    if (!seg) return fail("epde_batch_solve: no segment Grams (call epde_batch_gram first)");
    if (c->d_win.ensure(sizeof(double) * (size_t)b->win_doubles)) return 1;
    for (auto& g : b->groups) {
        WindowParams wp;
        wp.g = c->geom; wp.seg = seg + g.seg_off; wp.win = (double*)c->d_win.ptr + g.win_off; wp.n_ent = g.n_ent;
        wp.total = b->weighted ? seg + g.seg_off + (long long)c->geom.n_rows * g.n_ent : nullptr;
        KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
        window_gram_kernel<<<dim3(c->geom.n_windows + 1, (unsigned)((g.n_ent + 127) / 128)), 128, 0, c->stream>>>(wp);
        CK(cudaGetLastError());
    }
    if (b->n_eval > 0) {
        SolveParams sp;
        sp.g = c->geom; sp.eqs = b->d_eqs; sp.eval_eq = b->d_eval_eq; sp.eval_tgt = b->d_eval_tgt; sp.win = (const double*)c->d_win.ptr;
        sp.pairent = b->d_pairent;
        sp.mode = b->d_mode; sp.given_wterm = b->d_given_w; sp.given_b = b->d_given_b;
        sp.coef = b->d_coef; sp.intercept = b->d_intercept; sp.intercept_l1 = b->d_b1; sp.stability = b->d_stab; sp.sw = b->d_sw;
        sp.wterm = b->d_wterm; sp.n_active = b->d_nact; sp.all_zero = b->d_allzero; sp.iters = b->d_iters;
        sp.part = part; sp.nparts = nparts; sp.stage = 0;
        auto mine = [&](size_t n) { return (unsigned)((n > (size_t)part) ? (n - part + nparts - 1) / nparts : 0); };
        const size_t smem = sizeof(double) * (size_t)c->geom.n_windows * KS;
        if (mine(b->evals_gs8.size()) > 0) {
            sp.eval_list = b->d_lists + b->off_gs8;
            // the staged copy of the window Grams pays when the launch is latency-bound (few CTAs) or when it still
            // leaves room for three CTAs per SM; a large batch on a 3-D grid (W ~ 100) prefers the occupancy
            const size_t stage_bytes = sizeof(double) * (size_t)c->geom.n_windows * SOLVE_STAGE_DOUBLES;
            sp.stage = (mine(b->evals_gs8.size()) <= 2u * (unsigned)c->sm_count || smem + stage_bytes <= 40 * 1024) ? 1 : 0;
            { KernelSpan ks(c, EPDE_KERNEL_SOLVE); lasso_solve_kernel<8><<<mine(b->evals_gs8.size()), solve_threads(8), smem + (sp.stage ? stage_bytes : 0), c->stream>>>(sp); }
            sp.stage = 0;
            CK(cudaGetLastError());
        }
        if (mine(b->evals_gs16.size()) > 0) {
            sp.eval_list = b->d_lists + b->off_gs16;
            { KernelSpan ks(c, EPDE_KERNEL_SOLVE); lasso_solve_kernel<16><<<mine(b->evals_gs16.size()), SOLVE_THREADS, smem, c->stream>>>(sp); }
            CK(cudaGetLastError());
        }
        if (mine(b->evals_gs32.size()) > 0) {
            sp.eval_list = b->d_lists + b->off_gs32;
            { KernelSpan ks(c, EPDE_KERNEL_SOLVE); lasso_solve_kernel<32><<<mine(b->evals_gs32.size()), SOLVE_THREADS, smem, c->stream>>>(sp); }
            CK(cudaGetLastError());
        }
    }
    b->solved = true;
    return 0;
}

int epde_batch_solve(epde_batch* b, const double* dev_seg) {
    if (!b || !b->ctx) return fail("null batch");
    return solve_impl(b, dev_seg, 0, 1);
}

int epde_batch_solution_size(epde_batch* b, int with_sw, int64_t* n) {
    if (!b || !n) return fail("null argument");
    *n = (int64_t)((b->sol_bytes + (with_sw ? b->sw_bytes : 0)) / sizeof(double));
    return 0;
}

int epde_batch_solve_part(epde_batch* b, const double* dev_seg, int part, int nparts, int with_sw, double* dev_solution) {
    if (!b || !b->ctx || !dev_solution) return fail("epde_batch_solve_part: null argument");
    if (nparts < 1 || part < 0 || part >= nparts) return fail("epde_batch_solve_part: part %d of %d", part, nparts);
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    // evals of the other parts contribute zeros to the sum that merges the parts
    CK(cudaMemsetAsync(c->d_out.ptr, 0, b->sol_bytes, c->stream));
    if (with_sw) CK(cudaMemsetAsync(b->d_sw, 0, b->sw_bytes, c->stream));
    if (solve_impl(b, dev_seg, part, nparts)) return 1;
    CK(cudaMemcpyAsync(dev_solution, c->d_out.ptr, b->sol_bytes, cudaMemcpyDeviceToDevice, c->stream));
    if (with_sw)
        CK(cudaMemcpyAsync((char*)dev_solution + b->sol_bytes, b->d_sw, b->sw_bytes, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}

int epde_batch_set_solution(epde_batch* b, const double* dev_solution, int with_sw) {
    if (!b || !b->ctx || !dev_solution) return fail("epde_batch_set_solution: null argument");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(c->d_out.ptr, dev_solution, b->sol_bytes, cudaMemcpyDeviceToDevice, c->stream));
    if (with_sw)
        CK(cudaMemcpyAsync(b->d_sw, (const char*)dev_solution + b->sol_bytes, b->sw_bytes, cudaMemcpyDeviceToDevice, c->stream));
    b->solved = true;
    return 0;
}

int epde_batch_residual(epde_batch* b, double* dev_sums) {
    if (!b || !b->ctx) return fail("null batch");
    if (!b->solved) return fail("epde_batch_residual: call epde_batch_solve first");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    double* sums = dev_sums ? dev_sums : b->d_sums;
    if (b->n_eval == 0) return 0;
    const long long n_chunks = (c->N + PT_B - 1) / PT_B;
    // one persistent CTA per SM: the chunk -> CTA assignment (and with it the summation order) depends only on the
    // grid size, never on what else is in the batch
    const int grid = (int)std::min<long long>(n_chunks, (long long)c->sm_count);
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    for (auto& g : b->groups) {
        if (g.n_slots == 0) continue;
        const size_t n_pad = ((size_t)g.n_slots + OCT - 1) / OCT * OCT;
        const size_t o_col = align(sizeof(EvalJob) * n_pad), o_w = align(o_col + sizeof(int) * n_pad * TMAX),
                     o_hdr = align(o_w + sizeof(double) * n_pad * TMAX), o_ec = align(o_hdr + sizeof(EvalHdr) * n_pad),
                     o_ew = align(o_ec + sizeof(int) * n_pad * TMAX), o_meta = align(o_ew + sizeof(double) * n_pad * TMAX),
                     o_bl = o_meta + 64, tot = align(o_bl + sizeof(int) * (CMAX + 8));
        if (c->d_jobs.ensure(tot)) return 1;
        char* jb = (char*)c->d_jobs.ptr;
        ResidProgParams pp;
        pp.eqs = b->d_eqs; pp.eval_eq = b->d_eval_eq; pp.eval_tgt = b->d_eval_tgt; pp.slot_ev = g.d_slot_ev;
        pp.wterm = b->d_wterm; pp.intercept = b->d_intercept; pp.intercept_l1 = b->d_b1;
        pp.jobs = (EvalJob*)jb; pp.job_col = (int*)(jb + o_col); pp.job_w = (double*)(jb + o_w); pp.n_slots = g.n_slots;
        { KernelSpan ks(c, EPDE_KERNEL_RESIDUAL); residual_prog_kernel<<<(unsigned)((n_pad + 63) / 64), 64, 0, c->stream>>>(pp); }
        CK(cudaGetLastError());
        ResidPackParams sp;
        sp.jobs = pp.jobs; sp.job_col = pp.job_col; sp.job_w = pp.job_w; sp.n_slots = g.n_slots;
        sp.hdr = (EvalHdr*)(jb + o_hdr); sp.ent_col = (int*)(jb + o_ec); sp.ent_w = (double*)(jb + o_ew); sp.meta = (int*)(jb + o_meta);
        sp.blist = (int*)(jb + o_bl);
        sp.wcol = (b->weighted && b->want_l2) ? g.cp : -1;
        sp.nbuild = g.cp + (sp.wcol >= 0 ? 1 : 0);
        {
            int n2 = 1;
            while (n2 < g.n_slots) n2 <<= 1;
            KernelSpan ks(c, EPDE_KERNEL_RESIDUAL);
            residual_pack_kernel<<<1, 1024, n2 <= SORT_MAX ? (sizeof(unsigned long long) + sizeof(int)) * n2 : 0, c->stream>>>(sp);
        }
        CK(cudaGetLastError());
        ResidParams p;
        memset(&p, 0, sizeof p);
        p.N = c->N; p.U = (int)g.token_ids.size();
        for (int u = 0; u < p.U; ++u) p.tok[u] = c->tokens[g.token_ids[u]];
        p.cols = g.d_cols; p.gfac = g.d_gfac; p.nbuild = g.cp; p.wcol = -1;
        if (b->weighted && b->want_l2) { p.tok[p.U++] = c->weights; p.nbuild = g.cp + 1; p.wcol = g.cp; }
        p.hdr = sp.hdr; p.ent_col = sp.ent_col; p.ent_w = sp.ent_w; p.meta = sp.meta; p.blist = sp.blist;
        p.n_slots = g.n_slots; p.tcols = g.d_tcols; p.n_tcols = g.n_tcols;
        p.n_chunks = (int)n_chunks; p.aligned = 1;
        const int nsum = 2 * g.n_slots + g.n_tcols;
        if (c->d_partial.ensure(sizeof(double) * (size_t)grid * nsum)) return 1;
        p.partial = (double*)c->d_partial.ptr;
        const size_t tile_bytes = sizeof(double) * (size_t)p.U * PT_B;
        const size_t rest = sizeof(double) * ((size_t)p.nbuild * PP_B + nsum + 2) + sizeof(ColDesc) * (size_t)p.nbuild;
        p.nstages = (2 * tile_bytes + rest <= (size_t)c->smem_resid) ? 2 : 1;
        size_t smem = p.nstages * tile_bytes + rest + 16;
        if (smem > (size_t)c->smem_resid) return fail("pass B needs %zu bytes of shared memory (> %d)", smem, c->smem_resid);
        // whatever is left holds a copy of the eval descriptors (headers + an entry per surviving coefficient: at most
        // 32 + 12 * TMAX bytes per eval, typically ~50); the kernel falls back to global reads when they do not fit
        p.desc_smem = (int)std::min<size_t>((size_t)c->smem_resid - smem, n_pad * (sizeof(EvalHdr) + 12 * (size_t)TMAX) + 64);
        smem += (size_t)p.desc_smem;
        {
            KernelSpan ks(c, EPDE_KERNEL_RESIDUAL);
            if (b->want_l2) residual_cols_kernel<true><<<grid, RES_WARPS * 32, smem, c->stream>>>(p);
            else residual_cols_kernel<false><<<grid, RES_WARPS * 32, smem, c->stream>>>(p);
        }
        CK(cudaGetLastError());
        ResidFinalParams fp;
        fp.slot_ev = g.d_slot_ev; fp.slot_tci = g.d_slot_tci; fp.partial = p.partial; fp.sums = sums;
        fp.n_slots = g.n_slots; fp.n_tcols = g.n_tcols; fp.n_cta = grid; fp.want_l2 = b->want_l2 ? 1 : 0;
        { KernelSpan ks(c, EPDE_KERNEL_RESIDUAL); residual_final_kernel<<<(g.n_slots + 127) / 128, 128, 0, c->stream>>>(fp); }
        CK(cudaGetLastError());
    }
    return 0;
}

int epde_batch_fetch(epde_batch* b, const double* dev_sums, double* coef, double* intercept,
                     double* fitness_l1, double* fitness_l2, double* stability, double* sw_weights,
                     int32_t* n_active, int32_t* all_zero, int32_t* iters) {
    if (!b || !b->ctx) return fail("null batch");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    const int ne = b->n_eval;
    if (ne == 0) { CK(cudaStreamSynchronize(c->stream)); return 0; }
    if (dev_sums && dev_sums != b->d_sums)
        CK(cudaMemcpyAsync(b->d_sums, dev_sums, sizeof(double) * 3 * ne, cudaMemcpyDeviceToDevice, c->stream));
    // the pinned buffer doubles as the encoding staging area; the encoding copy is complete once
    // the stream reaches this point, so synchronise before reallocating / overwriting it
    CK(cudaStreamSynchronize(c->stream));
    const int W = c->geom.n_windows;
    const size_t sw_bytes = sw_weights ? sizeof(double) * (size_t)ne * W * KS : 0;
    if (ensure_pinned(c, b->out_bytes + sw_bytes + 256)) return 1;
    CK(cudaMemcpyAsync(c->h_pinned, c->d_out.ptr, b->out_bytes, cudaMemcpyDeviceToHost, c->stream));
    // the sliding-window weights (ne x W x KS doubles) are copied only when the caller wants them
    if (sw_weights) CK(cudaMemcpyAsync((char*)c->h_pinned + b->out_bytes, b->d_sw, sw_bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    const char* h = (const char*)c->h_pinned;
    const char* d0 = (const char*)c->d_out.ptr;
    auto hp = [&](const void* dptr) { return h + ((const char*)dptr - d0); };
    if (coef) memcpy(coef, hp(b->d_coef), sizeof(double) * ne * TMAX);
    if (intercept) memcpy(intercept, hp(b->d_intercept), sizeof(double) * ne);
    if (stability) memcpy(stability, hp(b->d_stab), sizeof(double) * ne);
    if (sw_weights) memcpy(sw_weights, h + b->out_bytes, sw_bytes);
    const double* hn = (const double*)hp(b->d_nact), *hz = (const double*)hp(b->d_allzero), *hi = (const double*)hp(b->d_iters);
    for (int v = 0; v < ne; ++v) {
        if (n_active) n_active[v] = (int32_t)hn[v];
        if (all_zero) all_zero[v] = (int32_t)hz[v];
        if (iters) iters[v] = (int32_t)hi[v];
    }
    const double* sums = (const double*)hp(b->d_sums);
    for (int v = 0; v < ne; ++v) {
        if (fitness_l1) fitness_l1[v] = sums[3 * v] / sums[3 * v + 1];
        if (fitness_l2) fitness_l2[v] = b->want_l2 ? std::sqrt(sums[3 * v + 2]) : std::nan("");
    }
    return 0;
}

int epde_evaluate_batch(epde_ctx* c, int n_eq, const int32_t* term_off, const int32_t* fac_off,
                        const int32_t* fac_token, const double* fac_power, const int32_t* fac_kind,
                        const double* fac_param, int n_eval, const int32_t* eval_eq,
                        const int32_t* eval_target, double* coef, double* intercept,
                        double* fitness_l1, double* fitness_l2, double* stability, double* sw_weights,
                        int32_t* n_active, int32_t* all_zero, int32_t* iters) {
    epde_batch* b = nullptr;
    if (epde_batch_create(c, n_eq, term_off, fac_off, fac_token, fac_power, fac_kind, fac_param,
                          n_eval, eval_eq, eval_target, &b)) return 1;
    int rc = epde_batch_gram(b, nullptr);
    if (!rc) rc = epde_batch_solve(b, nullptr);
    if (!rc) rc = epde_batch_residual(b, nullptr);
    if (!rc) rc = epde_batch_fetch(b, nullptr, coef, intercept, fitness_l1, fitness_l2, stability,
                                   sw_weights, n_active, all_zero, iters);
    std::string keep = g_err;
    epde_batch_destroy(b);
    if (rc) g_err = keep;
    return rc;
}
int epde_ctx_synchronize(epde_ctx* c) {
    if (!c) return fail("null context");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int epde_fd_gradient(epde_ctx* c, const double* dev_in, double* dev_out, int ndim, const int64_t* shape, int axis,
                     const double* coords) {
    if (!c || !dev_in || !dev_out || !shape || !coords) return fail("epde_fd_gradient: null argument");
    if (ndim < 1 || ndim > EPDE_MAX_NDIM || axis < 0 || axis >= ndim) return fail("epde_fd_gradient: bad ndim / axis");
    if (dev_in == dev_out) return fail("epde_fd_gradient: in-place differentiation is not supported");
    const int n = (int)shape[axis];
    if (n < 2) return fail("epde_fd_gradient: axis %d has %d points (numpy.gradient needs >= 2)", axis, n);
    CK(cudaSetDevice(c->device));
    FdParams p;
    memset(&p, 0, sizeof p);
    p.in = dev_in; p.out = dev_out; p.n = n;
    p.total = 1; p.stride = 1;
    for (int d = 0; d < ndim; ++d) { p.total *= shape[d]; if (d > axis) p.stride *= shape[d]; }
    std::vector<double> dx(n - 1);
    for (int i = 0; i + 1 < n; ++i) dx[i] = coords[i + 1] - coords[i];
    bool uniform = true;
    for (int i = 1; i + 1 < n; ++i) uniform = uniform && (dx[i] == dx[0]);
    p.uniform = uniform ? 1 : 0;
    p.dx = dx[0]; p.dx_first = dx[0]; p.dx_last = dx[n - 2];
    if (!uniform) {
        // numpy: a = -(dx2) / (dx1 * (dx1 + dx2)); b = (dx2 - dx1) / (dx1 * dx2); c = dx1 / (dx2 * (dx1 + dx2))
        std::vector<double> co(3 * (size_t)n, 0.0);
        for (int i = 1; i + 1 < n; ++i) {
            const volatile double dx1 = dx[i - 1], dx2 = dx[i];
            const volatile double s12 = dx1 + dx2;
            const volatile double d1 = dx1 * s12, d2 = dx1 * dx2, d3 = dx2 * s12;
            const volatile double diff = dx2 - dx1;
            co[i] = -(dx2) / d1; co[n + i] = diff / d2; co[2 * (size_t)n + i] = dx1 / d3;
        }
        if (c->d_fd.ensure(sizeof(double) * co.size())) return 1;
        CK(cudaMemcpyAsync(c->d_fd.ptr, co.data(), sizeof(double) * co.size(), cudaMemcpyHostToDevice, c->stream));
        CK(cudaStreamSynchronize(c->stream));      // `co` is a pageable temporary
        p.a = (const double*)c->d_fd.ptr; p.b = p.a + n; p.c = p.a + 2 * (size_t)n;
    }
    const int grid = (int)std::min<long long>((p.total + 255) / 256, (long long)c->sm_count * 16);
    c->launches++;
    if (p.total < (1ll << 31)) fd_gradient_kernel<unsigned><<<grid, 256, 0, c->stream>>>(p);
    else fd_gradient_kernel<long long><<<grid, 256, 0, c->stream>>>(p);
    CK(cudaGetLastError());
    return 0;
}

int epde_poly_derivative(epde_ctx* c, const double* dev_in, double* dev_out, int ndim, const int64_t* shape, int axis,
                         int order, int window, const double* weights, const double* coords) {
    if (!c || !dev_in || !dev_out || !shape || !weights || !coords) return fail("epde_poly_derivative: null argument");
    if (ndim < 1 || ndim > EPDE_MAX_NDIM || axis < 0 || axis >= ndim) return fail("epde_poly_derivative: bad ndim / axis");
    if (dev_in == dev_out) return fail("epde_poly_derivative: in-place differentiation is not supported");
    if (order < 1 || order > 3) return fail("epde_poly_derivative: derivative order %d not in 1..3 (fin_der.py:33)", order);
    if (window < 3 || !(window & 1)) return fail("epde_poly_derivative: window %d must be odd and >= 3", window);
    const int n = (int)shape[axis];
    const int bound = window / 2 + 1;                   // deriv_calculators.py:117
    if (n < 2 * bound + 1 || n < 6) return fail("epde_poly_derivative: axis %d has %d points, too few for window %d", axis, n, window);
    CK(cudaSetDevice(c->device));
    PolyParams p;
    memset(&p, 0, sizeof p);
    p.in = dev_in; p.out = dev_out; p.n = n; p.window = window; p.bound = bound; p.order = order;
    p.total = 1; p.stride = 1;
    for (int d = 0; d < ndim; ++d) { p.total *= shape[d]; if (d > axis) p.stride *= shape[d]; }
    const size_t wbytes = sizeof(double) * (size_t)n * window, xbytes = sizeof(double) * (size_t)n;
    if (c->d_fd.ensure(wbytes + xbytes)) return 1;
    CK(cudaMemcpyAsync(c->d_fd.ptr, weights, wbytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync((char*)c->d_fd.ptr + wbytes, coords, xbytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));               // pageable host sources
    p.w = (const double*)c->d_fd.ptr; p.x = (const double*)((char*)c->d_fd.ptr + wbytes);
    const int grid = (int)std::min<long long>((p.total + 255) / 256, (long long)c->sm_count * 16);
    c->launches++;
    if (p.total < (1ll << 31)) poly_derivative_kernel<unsigned><<<grid, 256, 0, c->stream>>>(p);
    else poly_derivative_kernel<long long><<<grid, 256, 0, c->stream>>>(p);
    CK(cudaGetLastError());
    return 0;
}

}  // extern "C"

// Scratch of the allocation-free entry point below: one fixed array per GPU that loads the library (held until
// exit).  In host code `g_noalloc_scratch` is only a stand-in for the symbol: its DEVICE address is looked up with
// cudaGetSymbolAddress, once per device with that device current, and that pointer is what the kernel is given.
static __device__ __align__(256) double g_noalloc_scratch[EPDE_NOALLOC_MAX_DOUBLES];
static std::mutex g_noalloc_mutex;
static double* g_noalloc_ptr[64] = {nullptr};

static int noalloc_scratch(int device, double** out) {          // the caller has made `device` current
    if (device < 0 || device >= 64) return fail("epde_poly_derivative_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail("epde_poly_derivative_noalloc: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = (double*)ptr;
    }
    *out = g_noalloc_ptr[device];
    return 0;
}

extern "C" {

int epde_poly_derivative_noalloc(epde_ctx* c, const double* dev_in, double* dev_out, int ndim, const int64_t* shape,
                                 int axis, int order, int window, const double* weights, const double* coords) {
    if (!c || !dev_in || !dev_out || !shape || !weights || !coords) return fail("epde_poly_derivative_noalloc: null argument");
    if (ndim < 1 || ndim > EPDE_MAX_NDIM || axis < 0 || axis >= ndim) return fail("epde_poly_derivative_noalloc: bad ndim / axis");
    if (dev_in == dev_out) return fail("epde_poly_derivative_noalloc: in-place differentiation is not supported");
    if (order < 1 || order > 3) return fail("epde_poly_derivative_noalloc: derivative order %d not in 1..3 (fin_der.py:33)", order);
    if (window < 3 || !(window & 1)) return fail("epde_poly_derivative_noalloc: window %d must be odd and >= 3", window);
    if (shape[axis] < 1 || shape[axis] > EPDE_NOALLOC_MAX_DOUBLES)
        return fail("epde_poly_derivative_noalloc: axis %d has %lld points; the fixed scratch holds %d doubles", axis,
                    (long long)shape[axis], EPDE_NOALLOC_MAX_DOUBLES);
    const int n = (int)shape[axis];
    const int bound = window / 2 + 1;                   // deriv_calculators.py:117
    if (n < 2 * bound + 1 || n < 6) return fail("epde_poly_derivative_noalloc: axis %d has %d points, too few for window %d", axis, n, window);
    const long long need = (long long)n * window + n;   // weights [n][window] + coordinates [n]
    if (need > EPDE_NOALLOC_MAX_DOUBLES)
        return fail("epde_poly_derivative_noalloc: %d points x (window %d + 1) = %lld doubles of scratch; the fixed scratch holds "
                    "%d (use epde_poly_derivative)", n, window, need, EPDE_NOALLOC_MAX_DOUBLES);
    CK(cudaSetDevice(c->device));
    double* scratch = nullptr;
    if (noalloc_scratch(c->device, &scratch)) return 1;
    PolyParams p;
    memset(&p, 0, sizeof p);
    p.in = dev_in; p.out = dev_out; p.n = n; p.window = window; p.bound = bound; p.order = order;
    p.total = 1; p.stride = 1;
    for (int d = 0; d < ndim; ++d) { p.total *= shape[d]; if (d > axis) p.stride *= shape[d]; }
    const size_t wbytes = sizeof(double) * (size_t)n * window, xbytes = sizeof(double) * (size_t)n;
    // every scratch value the kernel reads is overwritten here, so nothing of the previous call survives into this one
    CK(cudaMemcpyAsync(scratch, weights, wbytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(scratch + (size_t)n * window, coords, xbytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));               // pageable host sources
    p.w = scratch; p.x = scratch + (size_t)n * window;
    const int grid = (int)std::min<long long>((p.total + 255) / 256, (long long)c->sm_count * 16);
    c->launches++;
    if (p.total < (1ll << 31)) poly_derivative_kernel<unsigned><<<grid, 256, 0, c->stream>>>(p);
    else poly_derivative_kernel<long long><<<grid, 256, 0, c->stream>>>(p);
    CK(cudaGetLastError());
    return 0;
}

int epde_cache_upload_interior(epde_ctx* c, int id, const double* dev_full, int ndim, const int64_t* full_shape,
                               const int64_t* boundary) {
    if (!c || !dev_full || !full_shape || !boundary) return fail("epde_cache_upload_interior: null argument");
    if (ndim != c->ndim) return fail("epde_cache_upload_interior: ndim %d, context has %d", ndim, c->ndim);
    if (id < 0 || id >= EPDE_MAX_TOKENS) return fail("epde_cache_upload_interior: token id %d out of range", id);
    CropParams p;
    memset(&p, 0, sizeof p);
    p.in = dev_full; p.ndim = ndim; p.n_inner = c->N;
    long long stride = 1;
    for (int d = ndim - 1; d >= 0; --d) {
        if (boundary[d] < 0 || full_shape[d] - 2 * boundary[d] != c->shape[d])
            return fail("epde_cache_upload_interior: axis %d: full %lld - 2 x boundary %lld != inner %lld", d,
                        (long long)full_shape[d], (long long)boundary[d], c->shape[d]);
        p.inner[d] = c->shape[d]; p.off[d] = boundary[d]; p.full_stride[d] = stride;
        stride *= full_shape[d];
    }
    CK(cudaSetDevice(c->device));
    double*& tok = c->tokens[id];
    if (token_alloc(c, &tok, "epde_cache_upload_interior")) { c->tokens.erase(id); return 1; }
    p.out = tok;
    const int grid = (int)std::min<long long>((c->N + 255) / 256, (long long)c->sm_count * 16);
    c->launches++;
    if (stride < (1ll << 31)) crop_interior_kernel<unsigned><<<grid, 256, 0, c->stream>>>(p);      // stride = full-grid size here
    else crop_interior_kernel<long long><<<grid, 256, 0, c->stream>>>(p);
    CK(cudaGetLastError());
    return 0;
}

int epde_measure_fp64_peak(int device, double seconds, double* dfma_tflops, double* dmma_tflops) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail("epde_measure_fp64_peak: no such CUDA device %d", device);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    const int grid = prop.multiProcessorCount * 4, iters = 4000;
    double* out = nullptr;
    CK(cudaMalloc((void**)&out, sizeof(double) * (size_t)grid * 256));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double best[2] = {0.0, 0.0};
    for (int which = 0; which < 2; ++which) {
        const double flop = which == 0 ? 2.0 * grid * 256.0 * 8 * iters : 2.0 * grid * 8.0 * 8 * iters * 256.0;
        double spent = 0.0;
        for (int rep = 0; rep < 1000 && (rep < 3 || spent < 0.5 * seconds * 1e3); ++rep) {
            CK(cudaEventRecord(e0));
            if (which == 0) fp64_dfma_kernel<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9);
            else fp64_dmma_kernel<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            spent += ms;
            if (rep >= 1) best[which] = std::max(best[which], flop / (ms * 1e-3) * 1e-12);
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    if (dfma_tflops) *dfma_tflops = best[0];
    if (dmma_tflops) *dmma_tflops = best[1];
    return 0;
}

}  // extern "C"
