// epde_b200.cu -- B200 (sm_100a) kernels + C ABI of the EPDE fitness-evaluation engine.
//
// Path (reference paths relative to the EPDE tree):
//   token cache            epde/cache/cache.py:141-524              -> HBM-resident token table
//   term assembly          structure/factor.py:218-269, evaluators.py:92-151,
//                          structure/main_structures.py:207-235,588-641
//                                                                    -> gram_tile_kernel (pass A)
//   calculate_weights      epde/supplementary.py:394-471             -> segment Grams, window_gram_kernel,
//                                                                       lasso_solve_kernel
//   PhysicsInformedLasso   operators/common/sparsity.py:20-150       -> lasso_solve_kernel
//   L2LRFitness/L2Fitness  operators/common/fitness.py:27-183        -> residual_tile_kernel (pass B)
//
// Data layout in HBM
//   token t            double[N]   N = prod(inner_shape), C order, last axis contiguous
//   cell Grams  C      per equation: [cellA][cellB][piece][NACC]   NACC = T(T+3)/2
//                      (upper triangle of the T x T term Gram, then the T column sums)
//   segment Grams S    per equation: [segment row][NE], NE = NACC + 1 (last = point count);
//                      segment rows of all axes stacked.  This is the buffer that is all-reduced
//                      under grid sharding.
//   window Grams  Wg   per equation: [W + 1][NE]: row w = sum of the segment rows of sliding window w
//                      (direct summation, no prefix differences: a window that holds small values
//                      after a region of large ones keeps full relative accuracy); row W = all points.
//
// Execution model of the two streaming passes: a CTA stages one tile of every token the launch
// group references in shared memory (cp.async), then each WARP evaluates a different equation on
// that tile (lane <-> point column), so a token tile is read from L2/HBM once per group instead of
// once per equation and all per-point operand fetches are shared-memory loads.
//
// No CPU fallback exists in this file: every entry point either runs the CUDA path or fails.
#include "../../include/epde_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

#define TMAX EPDE_MAX_TERMS
#define FMAXF EPDE_MAX_FACTORS
#define KS (TMAX + 1)          // stride of a sliding-window weight row
#define RIDGE 1e-6             // supplementary.py:458
#define LASSO_MAX_ITER 20      // sparsity.py:21
#define LASSO_TOL 1e-4         // sparsity.py:21

#define TILE_ROWS 64           // rows (leading-axes index pairs) of a pass-A tile
#define TILE_COLS 32           // columns (last axis) of a pass-A tile = warp width
#define GRAM_WARPS 8           // warps per pass-A CTA
#define MAXU 12                // distinct tokens per launch group (tile = MAXU x 64 x 32 doubles = 192 KB)
#define RES_WARPS 8            // warps per pass-B CTA (8 x <= 255 registers; 12 and 16 warps measured slower)
#define RES_TILE_MIN 1024      // points of a pass-B tile

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
struct FactorDesc {          // 32 bytes
    int slot;                // token slot inside the launch group's tile
    int ipow;                // integer power if power is 1, 2 or 3, else 0
    int kind;                // 0 pow, 1 sin, 2 cos, 3 inverse
    int pad;
    double power;
    double param;            // freq for trig kinds
};

struct EqDesc {
    int T;
    int fast;                // 1: every term has <= 2 factors, all kind 0 with power in {1, 2, 3}
    int group;               // launch group
    int fac_off[TMAX + 1];   // into the batch factor array
    long long s_off;         // doubles, into the segment-Gram table
    long long c_off;         // doubles, into the cell-Gram scratch of its launch group
    long long w_off;         // doubles, into the window-Gram table
    int4 fastop[TMAX];       // fast terms IN EVALUATION ORDER (most slot factors first): slots {A, C, B, D} of
                             // the term's factors with integer powers expanded (u^2 * v -> u, u, v); -1 = absent
    int2 fastop2[TMAX];      // slots {E, F} of terms with five / six slot factors
    int n2, n3;              // terms with >= 2 slot factors; depth: 0 (<= 2 factors everywhere), 1 (<= 4), 2 (<= 6)
    unsigned char pos[TMAX]; // term index -> evaluation position
    unsigned char perm[TMAX];// evaluation position -> term index
    unsigned char accmap[TMAX * (TMAX + 3) / 2];   // Gram accumulator in evaluation order -> index in term order
};

struct UnitDesc {            // one CTA work unit of pass A: a (sub-)cell of the leading axes x a column chunk
    int a_lo, a_hi;          // index range on leading slot A
    int b_lo, b_hi;          // index range on leading slot B
    int col_lo, ncols;       // columns (last axis) handled, ncols <= TILE_COLS
    int cell;                // (cellA * nB + cellB)
    int piece_lo, npieces;   // pieces (sub-segments of the last axis) covered by the chunk
    int wlog;                // >= 0: every piece of the unit is exactly 2^wlog columns wide (shuffle flush); -1: general
    int pad1, pad2;
};

struct MargDesc {            // one output row of the segment-Gram table
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
    int aligned16;                 // 1: every tile row starts on a 16-byte boundary
    long long N;
    const UnitDesc* units;
    const int* piece_start;        // [npieces + 1] column bounds
    const MargDesc* marg;          // [n_rows]
    const int* win_lo_row;         // [W + 1] first segment row of window w (entry W: all of axis 0)
    const int* win_hi_row;         // [W + 1] one past the last segment row
};

__host__ __device__ __forceinline__ int nacc_of(int T) { return T * (T + 3) / 2; }

// index of the (i, j) product inside a Gram row; column T is the ones column
__host__ __device__ __forceinline__ int gidx(int i, int j, int T) {
    if (i > j) { int t = i; i = j; j = t; }
    if (j == T) return (i == T) ? nacc_of(T) : T * (T + 1) / 2 + i;
    return i * T - i * (i - 1) / 2 + (j - i);
}

// ------------------------------------------------------------------------------------------
// shared-memory staging helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
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
    return pow(x, f.power);
}

__device__ __forceinline__ double ipow123(double x, int ip) {
    if (ip >= 2) { const double s = x * x; x = (ip == 3) ? s * x : s; }
    return x;
}

// generic term: product of its factors; `pt` = tile base + point offset, slot stride `tstride`
__device__ __forceinline__ double eval_term_generic(const FactorDesc* fac, int f0, int f1, const double* pt,
                                                    int tstride) {
    double v = 1.0;
    for (int f = f0; f < f1; ++f) {
        const FactorDesc d = fac[f];
        v *= apply_factor(d, pt[d.slot * tstride]);
    }
    return v;
}

// per-warp copy of the equation being processed (shared memory)
struct WarpProg {
    int4 op[TMAX];           // fast terms in evaluation order: byte offsets of slots {A, C, B, D}, -1 = absent
    int foff[TMAX + 1];      // generic path: factor ranges
    int2 op2[TMAX];          // byte offsets of slots {E, F}
    int T, fast, n2, n3;
    long long c_off;         // doubles, into the cell-Gram scratch of the launch group
    unsigned char accmap[TMAX * (TMAX + 3) / 2 + 4];
};

// ------------------------------------------------------------------------------------------
// pass A: fused term assembly + cell Gram.
// Persistent CTAs (one per SM) walk the units (a (sub-)cell of the leading axes x <= 32 columns of the
// last axis): the tile of every token of the launch group is staged in shared memory once, then each
// warp takes equations from the group's list (sorted by code variant, dynamic) and accumulates that
// equation's cell Gram in registers, lane <-> column, looping over the tile rows.  The per-lane
// accumulators are reduced over the columns of each last-axis piece with warp shuffles only (xor
// butterfly when the pieces have one power-of-two width, segmented shfl_down otherwise) and written
// to the cell table (no atomics: fixed summation order).
// ------------------------------------------------------------------------------------------
struct GramParams {
    Geom g;
    const EqDesc* eqs;
    const FactorDesc* factors;
    const int* eq_list;      // equations of this launch group, most terms first
    int n_eq;
    int U;                   // tokens of the group
    const double* tok[MAXU];
    double* cells;           // cell-Gram scratch of the group
};

// explicit shared-space accesses: the per-equation worker functions are not inlined, so a generic
// pointer would cost 64-bit address arithmetic and generic loads on every operand fetch
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ double lds_f64(unsigned a) {
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ int4 lds_v4(unsigned a) {
    int4 v;
    asm("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ double lds_f64_volatile(unsigned a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f64_volatile(unsigned a, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory");
}

template <int T>
__device__ __forceinline__ void gram_accumulate(double (&acc)[T * (T + 3) / 2], const double (&c)[T]) {
    int k = 0;
#pragma unroll
    for (int i = 0; i < T; ++i)
#pragma unroll
        for (int j = i; j < T; ++j) { acc[k] = fma(c[i], c[j], acc[k]); ++k; }
#pragma unroll
    for (int i = 0; i < T; ++i) { acc[k] += c[i]; ++k; }
}

// fast terms for R consecutive tile rows of one lane, branch free.  A fast term is a product of 1..4 token
// values (integer powers expanded: u^2 * v = u * u * v).  Terms come in evaluation order, most factors
// first.  N2 = number of leading terms that get a second-factor multiply, one of {0, 1, 2, T} (the class
// >= the equation's count); DEEP 1 / 2: those N2 terms also get factors three-four / three-six.  A term inside a
// class without that factor multiplies by 1.0 read from a small shared region of ones, so the body is
// straight-line code the scheduler interleaves with the Gram FMAs.  ops: shared address of the int4 op
// table (byte offsets of slots {A, C, B, D}, -1 = absent); pr: shared address of the lane's element in
// row `row` of slot 0; rows are ROWB bytes apart; ones: shared address of >= R * ROWB bytes of 1.0.
template <int T, int N2, int DEEP>
struct TermOffsets {
    int a[T];
    int b[N2 > 0 ? N2 : 1];
    int c[(DEEP >= 1 && N2 > 0) ? N2 : 1], d[(DEEP >= 1 && N2 > 0) ? N2 : 1];
    int e[(DEEP >= 2 && N2 > 0) ? N2 : 1], f[(DEEP >= 2 && N2 > 0) ? N2 : 1];
};

template <int T, int N2, int DEEP>
__device__ __forceinline__ void load_term_offsets(unsigned ops, unsigned ops2, TermOffsets<T, N2, DEEP>& o) {
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const int4 q = lds_v4(ops + t * 16);
        o.a[t] = q.x;
        if (t < N2) {
            o.b[t] = q.z;
            if (DEEP >= 1) { o.c[t] = q.y; o.d[t] = q.w; }
            if (DEEP >= 2) {
                int ex, fx;
                asm("ld.shared.v2.s32 {%0, %1}, [%2];" : "=r"(ex), "=r"(fx) : "r"(ops2 + t * 8));
                o.e[t] = ex; o.f[t] = fx;
            }
        }
    }
}

template <int T, int N2, int DEEP, int R, int ROWB>
__device__ __forceinline__ void fast_terms_r(const TermOffsets<T, N2, DEEP>& o, unsigned pr, unsigned ones,
                                             double (&c)[R][T]) {
#pragma unroll
    for (int t = 0; t < T; ++t) {
#pragma unroll
        for (int r = 0; r < R; ++r) c[r][t] = lds_f64(pr + o.a[t] + r * ROWB);
    }
#pragma unroll
    for (int t = 0; t < N2; ++t) {
        const unsigned ab = (o.b[t] >= 0) ? pr + o.b[t] : ones;
#pragma unroll
        for (int r = 0; r < R; ++r) c[r][t] *= lds_f64(ab + r * ROWB);
    }
    if (DEEP >= 1) {
#pragma unroll
        for (int t = 0; t < N2; ++t) {
            const unsigned ac = (o.c[t] >= 0) ? pr + o.c[t] : ones;
            const unsigned ad = (o.d[t] >= 0) ? pr + o.d[t] : ones;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                c[r][t] *= lds_f64(ac + r * ROWB);
                c[r][t] *= lds_f64(ad + r * ROWB);
            }
        }
    }
    if (DEEP >= 2) {
#pragma unroll
        for (int t = 0; t < N2; ++t) {
            const unsigned ae = (o.e[t] >= 0) ? pr + o.e[t] : ones;
            const unsigned af = (o.f[t] >= 0) ? pr + o.f[t] : ones;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                c[r][t] *= lds_f64(ae + r * ROWB);
                c[r][t] *= lds_f64(af + r * ROWB);
            }
        }
    }
}

// class of a two-factor count: the smallest of {0, 1, 2, T} that is >= n2, as an index 0..3.
// coarse (pass B): N2 in {0, 2, T} and DEPTH in {0, 2} only -- half the code variants, so fewer instruction-cache
// misses when the warps of a CTA run different equations, at the price of a few more multiplies by 1.0
// (measured at 512^3: pass B 128 -> 122 ms; pass A, which is shared-memory-bandwidth bound, 53 -> 58 ms, so
// pass A keeps the fine classes)
#define COARSE_CLASSES_B 0
__device__ __forceinline__ int n2_class(int n2, bool coarse) {
    if (coarse) return n2 == 0 ? 0 : (n2 <= 2 ? 2 : 3);
    return n2 <= 2 ? n2 : 3;
}

// flush: sum the lanes (columns) of every last-axis piece of the unit with warp shuffles (fixed tree order, no
// shared memory) and let the first lane of each piece store its NACC sums, un-permuted through accmap.
//   wlog >= 0: every piece is 2^wlog lanes wide and starts on a multiple of it -> xor butterfly in the lane groups;
//   otherwise: segmented reduction by doubling; lane_piece = piece of this lane's column (-1 beyond the unit).
template <int T>
__device__ __forceinline__ void flush_cells(const double (&acc)[T * (T + 3) / 2], int lane, double* out, int npieces,
                                            const unsigned char* accmap, int wlog, int lane_piece) {
    constexpr int NACC = T * (T + 3) / 2;
    double v[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) v[k] = acc[k];
    bool head;
    int pp;
    if (wlog >= 0) {
        for (int off = 1; off < (1 << wlog); off <<= 1) {
#pragma unroll
            for (int k = 0; k < NACC; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
        }
        pp = lane >> wlog;
        head = (lane & ((1 << wlog) - 1)) == 0 && pp < npieces;
    } else {
        for (int off = 1; off < 32; off <<= 1) {
            const int other = __shfl_down_sync(0xffffffffu, lane_piece, off);
            const bool take = (lane + off < 32) && (other == lane_piece);
#pragma unroll
            for (int k = 0; k < NACC; ++k) {
                const double t = __shfl_down_sync(0xffffffffu, v[k], off);
                if (take) v[k] += t;
            }
        }
        const int prev = __shfl_up_sync(0xffffffffu, lane_piece, 1);
        pp = lane_piece;
        head = pp >= 0 && (lane == 0 || prev != lane_piece);
    }
    if (head) {
        double* o = out + (long long)pp * NACC;
#pragma unroll
        for (int k = 0; k < NACC; ++k) o[accmap[k]] = v[k];
    }
    __syncwarp();
}

// fast path: every term = x^p [* y^q], p, q in {1, 2, 3}; R tile rows per iteration
template <int T, int N2, int DEEP, int R>
__device__ __noinline__ void gram_warp_fast(unsigned tile, unsigned ops, unsigned ops2, unsigned ones, int nrows, int lane,
                                            double* out, int npieces, const unsigned char* accmap, int wlog,
                                            int lane_piece) {
    constexpr int NACC = T * (T + 3) / 2;
    constexpr int ROWB = TILE_COLS * 8;
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
    const unsigned base = tile + lane * 8;
    TermOffsets<T, N2, DEEP> off;
    load_term_offsets<T, N2, DEEP>(ops, ops2, off);
    int row = 0;
    if (R > 1) {
        for (; row + R <= nrows; row += R) {
            double c[R][T];
            fast_terms_r<T, N2, DEEP, R, ROWB>(off, base + row * ROWB, ones, c);
#pragma unroll
            for (int r = 0; r < R; ++r) gram_accumulate<T>(acc, c[r]);
        }
    }
    for (; row < nrows; ++row) {
        double c[1][T];
        fast_terms_r<T, N2, DEEP, 1, ROWB>(off, base + row * ROWB, ones, c);
        gram_accumulate<T>(acc, c[0]);
    }
    flush_cells<T>(acc, lane, out, npieces, accmap, wlog, lane_piece);
}

// generic path: any factor kind / power / factor count
template <int T>
__device__ __noinline__ void gram_warp_generic(const double* tile, int tstride, const WarpProg* wp,
                                               const FactorDesc* fac, int nrows, int lane, double* out, int npieces,
                                               int wlog, int lane_piece) {
    constexpr int NACC = T * (T + 3) / 2;
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
    const double* base = tile + lane;
    for (int row = 0; row < nrows; ++row) {
        double c[T];
#pragma unroll
        for (int t = 0; t < T; ++t)
            c[t] = eval_term_generic(fac, wp->foff[t], wp->foff[t + 1], base + row * TILE_COLS, tstride);
        gram_accumulate<T>(acc, c);
    }
    flush_cells<T>(acc, lane, out, npieces, wp->accmap, wlog, lane_piece);
}

template <int T>
__device__ __forceinline__ void gram_warp(const double* tile, int tstride, const WarpProg* wp, const FactorDesc* fac,
                                          unsigned ones, int nrows, int lane, double* out, int npieces, int wlog,
                                          int lane_piece) {
    // 8-warp CTAs may use 255 registers per thread: four rows per iteration (two for the widest equations)
    constexpr int R = T <= 6 ? 4 : 2;
    if (wp->fast) {
#define GWF(NN, DD) gram_warp_fast<T, (NN), (DD), R>(smem_addr(tile), smem_addr(wp->op), smem_addr(wp->op2), ones, nrows, \
                                                     lane, out, npieces, wp->accmap, wlog, lane_piece)
#define GWD(NN) do { if (wp->n3 == 0) GWF(NN, 0); else if (wp->n3 == 1) GWF(NN, 1); else GWF(NN, 2); } while (0)
        switch (n2_class(wp->n2, false)) {
            case 0: GWF(0, 0); break;
            case 1: GWD(1); break;
            case 2: if (T >= 2) { GWD(T >= 2 ? 2 : T); break; }
            default: GWD(T); break;
        }
#undef GWD
#undef GWF
    } else {
        gram_warp_generic<T>(tile, tstride, wp, fac, nrows, lane, out, npieces, wlog, lane_piece);
    }
}

template <int NW>
__global__ void __launch_bounds__(NW * 32, 1) gram_tile_kernel(GramParams p) {
    extern __shared__ __align__(16) double smem[];
    __shared__ int s_next;
    __shared__ const double* s_tok[MAXU];
    __shared__ __align__(16) double s_ones[128];     // 4 rows x 256 bytes of 1.0 (fast_terms_r)
    __shared__ int s_lane_piece[TILE_COLS];
    const int tid = threadIdx.x;
    const int nwarps = blockDim.x >> 5;
    const int warp = tid >> 5, lane = tid & 31;
    for (int q = tid; q < 128; q += blockDim.x) s_ones[q] = 1.0;
    const int tstride = p.g.max_rows * TILE_COLS;
    double* tile = smem;                                              // [U][max_rows][32]
    WarpProg* progs = (WarpProg*)(tile + (size_t)p.U * tstride);      // [n_eq]
    if (tid < MAXU) s_tok[tid] = p.tok[tid];

    // ---- per-equation programs, built once per (persistent) CTA: a task then starts from shared memory only
    for (int le = warp; le < p.n_eq; le += nwarps) {
        const EqDesc* eq = p.eqs + p.eq_list[le];
        WarpProg* wp = progs + le;
        if (lane < TMAX) {
            int4 d = eq->fastop[lane];            // slots -> byte offsets inside the tile
            d.x *= tstride * 8;
            if (d.y >= 0) d.y *= tstride * 8;
            if (d.z >= 0) d.z *= tstride * 8;
            if (d.w >= 0) d.w *= tstride * 8;
            wp->op[lane] = d;
            int2 d2 = eq->fastop2[lane];
            if (d2.x >= 0) d2.x *= tstride * 8;
            if (d2.y >= 0) d2.y *= tstride * 8;
            wp->op2[lane] = d2;
        }
        if (lane <= TMAX) wp->foff[lane] = eq->fac_off[lane];
        if (lane == 0) { wp->T = eq->T; wp->fast = eq->fast; wp->n2 = eq->n2; wp->n3 = eq->n3; wp->c_off = eq->c_off; }
        for (int k = lane; k < TMAX * (TMAX + 3) / 2; k += 32) wp->accmap[k] = eq->accmap[k];
    }
    __syncthreads();

    for (int ui = blockIdx.x; ui < p.g.n_units; ui += gridDim.x) {
        const UnitDesc u = p.g.units[ui];
        const int wb = u.b_hi - u.b_lo;
        const int nrows = (u.a_hi - u.a_lo) * wb;
        // ---- stage the tile: 16-byte cp.async where rows are aligned and full, 8-byte / zero fill otherwise
        {
            const bool vec = p.g.aligned16 && !(u.col_lo & 1);
            const int total = p.U * nrows * (TILE_COLS / 2);
            for (int q = tid; q < total; q += blockDim.x) {
                const int part = q & (TILE_COLS / 2 - 1);
                const int rr = q / (TILE_COLS / 2);
                const int row = rr % nrows, us = rr / nrows;
                const int ia = u.a_lo + row / wb, ib = u.b_lo + row % wb;
                const int c0 = part * 2;
                const double* src = s_tok[us] + ((long long)ia * p.g.strideA + (long long)ib * p.g.strideB + u.col_lo + c0);
                double* dst = tile + (size_t)us * tstride + row * TILE_COLS + c0;
                if (vec && c0 + 2 <= u.ncols) {
                    cp_async16(dst, src);
                } else {
                    if (c0 < u.ncols) cp_async8(dst, src); else dst[0] = 0.0;
                    if (c0 + 1 < u.ncols) cp_async8(dst + 1, src + 1); else dst[1] = 0.0;
                }
            }
            // piece of every tile column (general flush path)
            if (tid < TILE_COLS) {
                int pc = -1;
                for (int q = 0; q < u.npieces; ++q) {
                    const int lo = p.g.piece_start[u.piece_lo + q] - u.col_lo, hi = p.g.piece_start[u.piece_lo + q + 1] - u.col_lo;
                    if (tid >= lo && tid < hi) pc = q;
                }
                s_lane_piece[tid] = pc;
            }
            cp_async_wait_all();
            if (tid == 0) s_next = nwarps;
            __syncthreads();
        }
        const int lane_piece = s_lane_piece[lane];
        int le = warp;
        while (le < p.n_eq) {
            const WarpProg* wp = progs + le;
            const int T = wp->T;
            double* out = p.cells + wp->c_off + ((long long)u.cell * p.g.npieces + u.piece_lo) * nacc_of(T);
#define GW(TT) gram_warp<TT>(tile, tstride, wp, p.factors, smem_addr(s_ones), nrows, lane, out, u.npieces, u.wlog, lane_piece)
            switch (T) {
                case 1: GW(1); break;
                case 2: GW(2); break;
                case 3: GW(3); break;
                case 4: GW(4); break;
                case 5: GW(5); break;
                case 6: GW(6); break;
                case 7: GW(7); break;
                default: GW(8); break;
            }
#undef GW
            if (lane == 0) le = atomicAdd(&s_next, 1);
            le = __shfl_sync(0xffffffffu, le, 0);
        }
        __syncthreads();            // everyone is done with the tile before it is refilled
    }
}

// ------------------------------------------------------------------------------------------
// marginals: segment Gram row = sum of the cell Grams of the slab, in a fixed order (deterministic)
// grid = (n_rows, n_eq_group); block = 256: thread <-> (cell lane, Gram entry) with 256 / NACC cell lanes, so
// consecutive threads read consecutive doubles of consecutive cells (the slab of a leading-axis row is one
// contiguous range of the cell table) and four independent loads per thread are in flight.
// ------------------------------------------------------------------------------------------
struct MargParams {
    Geom g;
    const EqDesc* eqs;
    const int* eq_list;
    const double* cells;
    double* seg;
};

__global__ void __launch_bounds__(256) marginal_kernel(MargParams p) {
    __shared__ double part[256];
    const int e = p.eq_list[blockIdx.y];
    const EqDesc& eq = p.eqs[e];
    const int NACC = nacc_of(eq.T);
    const int NE = NACC + 1;
    const MargDesc m = p.g.marg[blockIdx.x];
    const int tid = threadIdx.x;
    const int cpt = 256 / NACC;                    // cells per trip (>= 5: NACC <= 44)
    const int sub = tid / NACC, k = tid - sub * NACC;
    const unsigned nb = m.b_hi - m.b_lo, np = m.p_hi - m.p_lo;
    const int ncell = (m.a_hi - m.a_lo) * (int)nb * (int)np;
    const double* C = p.cells + eq.c_off + k;
    auto at = [&](int q) -> double {
        const unsigned r = (unsigned)q / np, ip = (unsigned)q - r * np;
        const unsigned iaa = r / nb, ibb = r - iaa * nb;
        const long long cell = ((long long)(m.a_lo + (int)iaa) * p.g.nB + (m.b_lo + (int)ibb)) * p.g.npieces + (m.p_lo + (int)ip);
        return __ldcs(C + cell * NACC);             // read once: do not keep in L2 ahead of the token tiles
    };
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    if (sub < cpt) {
        int q = sub;
        for (; q + 3 * cpt < ncell; q += 4 * cpt) {
            const double v0 = at(q), v1 = at(q + cpt), v2 = at(q + 2 * cpt), v3 = at(q + 3 * cpt);
            s0 += v0; s1 += v1; s2 += v2; s3 += v3;
        }
        for (; q < ncell; q += cpt) s0 += at(q);
    }
    part[tid] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (tid < NACC) {
        double s = 0.0;
        for (int j = 0; j < cpt; ++j) s += part[j * NACC + tid];
        p.seg[eq.s_off + (long long)m.row * NE + tid] = s;
    }
    if (tid == 64) p.seg[eq.s_off + (long long)m.row * NE + NACC] = m.count;
}

// Three-axis grids: every cell belongs to one row of each axis, so the kernel above reads the cell table three
// times.  Two steps read it once: step 1, one CTA per (slot-A cell index a, equation), walks the contiguous slab
// [a][all b][all pieces] and leaves  tot[a], byB[a][b], byP[a][piece]  (byP in per-thread shared slots: thread
// (cell lane, entry) always meets the same pieces); step 2 sums those over the ranges of each segment row.
struct MargSlabParams {
    Geom g;
    const EqDesc* eqs;
    const int* eq_list;
    const double* cells;
    double* part;            // per equation of the group: [nA][1 + nB + npieces][NACC]
    long long eq_stride;
    double* seg;
};

__global__ void __launch_bounds__(256) marginal_slab_kernel(MargSlabParams p) {
    extern __shared__ double s_pp[];               // [npieces][NACC]
    __shared__ double part[256];
    const int a = blockIdx.x;
    const EqDesc& eq = p.eqs[p.eq_list[blockIdx.y]];
    const int NACC = nacc_of(eq.T);
    const int tid = threadIdx.x;
    const int cpt = 256 / NACC;
    const int sub = tid / NACC, k = tid - sub * NACC;
    const int nB = p.g.nB, np = p.g.npieces;
    const bool act = sub < cpt;
    if (act)
        for (int q = sub; q < np; q += cpt) s_pp[q * NACC + k] = 0.0;
    const double* C = p.cells + eq.c_off + (long long)a * nB * np * NACC + k;
    double* out = p.part + (long long)blockIdx.y * p.eq_stride + (long long)a * (1 + nB + np) * NACC;
    double tot = 0.0;
    for (int b = 0; b < nB; ++b) {
        double vb = 0.0;
        if (act) {
            const double* Cb = C + (long long)b * np * NACC;
#pragma unroll 4
            for (int q = sub; q < np; q += cpt) {
                const double v = __ldcs(Cb + q * NACC);
                vb += v;
                s_pp[q * NACC + k] += v;
            }
        }
        part[tid] = vb;
        __syncthreads();
        if (tid < NACC) {
            double t = 0.0;
            for (int j = 0; j < cpt; ++j) t += part[j * NACC + tid];
            out[(1 + b) * NACC + tid] = t;
            tot += t;
        }
        __syncthreads();
    }
    if (tid < NACC) out[tid] = tot;
    if (act)
        for (int q = sub; q < np; q += cpt) out[(1 + nB + q) * NACC + k] = s_pp[q * NACC + k];
}

// grid = (n_rows, n_eq_group), block = 64
__global__ void __launch_bounds__(64) marginal_rows_kernel(MargSlabParams p) {
    const EqDesc& eq = p.eqs[p.eq_list[blockIdx.y]];
    const int NACC = nacc_of(eq.T);
    const int NE = NACC + 1;
    const MargDesc m = p.g.marg[blockIdx.x];
    const int k = threadIdx.x;
    const int rowlen = 1 + p.g.nB + p.g.npieces;
    if (k < NACC) {
        const double* P = p.part + (long long)blockIdx.y * p.eq_stride + k;
        double s = 0.0;
        if (m.kind == 0) {
#pragma unroll 4
            for (int a = m.a_lo; a < m.a_hi; ++a) s += P[(long long)a * rowlen * NACC];
        } else {
            const int lo = (m.kind == 1) ? 1 + m.b_lo : 1 + p.g.nB + m.p_lo;
            const int n = (m.kind == 1) ? m.b_hi - m.b_lo : m.p_hi - m.p_lo;
            for (int q = 0; q < n; ++q) {
                double t0 = 0.0, t1 = 0.0;
                int a = 0;
                for (; a + 1 < p.g.nA; a += 2) {
                    t0 += P[((long long)a * rowlen + lo + q) * NACC];
                    t1 += P[((long long)(a + 1) * rowlen + lo + q) * NACC];
                }
                if (a < p.g.nA) t0 += P[((long long)a * rowlen + lo + q) * NACC];
                s += t0 + t1;
            }
        }
        p.seg[eq.s_off + (long long)m.row * NE + k] = s;
    }
    if (k == 63) p.seg[eq.s_off + (long long)m.row * NE + NACC] = m.count;
}

// window Gram = direct sum of the segment rows of the window; row W = every point (all of axis 0).
// grid = (W + 1, n_eq), block = 64
struct WindowParams {
    Geom g;
    const EqDesc* eqs;
    const double* seg;
    double* win;
};

__global__ void window_gram_kernel(WindowParams p) {
    const EqDesc& eq = p.eqs[blockIdx.y];
    const int NE = nacc_of(eq.T) + 1;
    const int w = blockIdx.x;
    const int r0 = p.g.win_lo_row[w], r1 = p.g.win_hi_row[w];
    const double* S = p.seg + eq.s_off;
    for (int a = threadIdx.x; a < NE; a += blockDim.x) {
        double s = 0.0;
        for (int r = r0; r < r1; ++r) s += S[(long long)r * NE + a];
        p.win[eq.w_off + (long long)w * NE + a] = s;
    }
}
// ------------------------------------------------------------------------------------------
// solve: one CTA per eval.  calculate_weights (supplementary.py:455-469) = every window's ridge
// system solved by a GS-lane group (Gaussian elimination with partial pivoting, as LAPACK gesv);
// PhysicsInformedLasso.fit (sparsity.py:42-150) on the total Gram by thread 0 with the reference's
// exact control flow.
// ------------------------------------------------------------------------------------------
struct SolveParams {
    Geom g;
    const EqDesc* eqs;
    const int* eval_list;
    const int* eval_eq;
    const int* eval_tgt;
    const double* win;       // window Grams [W + 1][NE] per equation (row W = all points)
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
    int* n_active;
    int* all_zero;
    int* iters;
};

#define SOLVE_THREADS 128

__device__ __forceinline__ double nan_to_num(double x) {
    if (isnan(x)) return 0.0;
    if (isinf(x)) return x > 0 ? 1.7976931348623157e308 : -1.7976931348623157e308;
    return x;
}

template <int GS>
__global__ void __launch_bounds__(SOLVE_THREADS) lasso_solve_kernel(SolveParams p) {
    constexpr int NG = SOLVE_THREADS / GS;
    __shared__ double Gt[KS * KS];             // total Gram of all points, symmetric K x K
    __shared__ double M[NG][GS][GS + 1];       // augmented systems
    __shared__ int cols[KS];                   // columns of the current system: active terms, then T (ones)
    __shared__ double s_coef[TMAX], s_cv[TMAX], s_mu[KS], s_cvraw[KS];
    __shared__ int s_n, s_state, s_iter, s_sweeps;
    __shared__ double s_b;
    extern __shared__ double s_sw[];           // [W][KS]

    const int tid = threadIdx.x;
    const int ev = p.eval_list[blockIdx.x];
    const int e = p.eval_eq[ev];
    const int tgt = p.eval_tgt[ev];
    const EqDesc& eq = p.eqs[e];
    const int T = eq.T, F = T - 1;
    const int NE = nacc_of(T) + 1;
    const int W = p.g.n_windows;
    const double* P = p.win + eq.w_off;

    // total Gram
    for (int q = tid; q < (T + 1) * (T + 1); q += SOLVE_THREADS) {
        const int i = q / (T + 1), j = q % (T + 1);
        Gt[i * KS + j] = P[(long long)W * NE + gidx(i, j, T)];
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
                for (int c = 0; c < n; ++c) {
                    const int gi = gidx(i, cols[c], T);
                    double v = Phi[gi];
                    if (c == lr) v += RIDGE;
                    Mg[lr][c] = v;
                }
                const int gi = gidx(i, tgt, T);
                Mg[lr][n] = Phi[gi];
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
                if (idx != k && lr >= k && lr <= n) {
                    const double t0 = Mg[k][lr];
                    Mg[k][lr] = Mg[idx][lr];
                    Mg[idx][lr] = t0;
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
        if (tid < n) {
            double s = 0.0;
            for (int w = 0; w < W; ++w) s += s_sw[w * KS + tid];
            const double mu = s / W;
            double q = 0.0;
            for (int w = 0; w < W; ++w) { const double dlt = s_sw[w * KS + tid] - mu; q = fma(dlt, dlt, q); }
            const double var = q / (W - 1);     // std(ddof=1)**2 ; W == 1 -> nan like numpy
            double cv = var / (mu * mu);
            if (mu == 0.0) cv = 0.0;
            s_mu[tid] = mu;
            s_cvraw[tid] = cv;
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
            auto all_zero = [&]() { for (int j = 0; j < F; ++j) if (s_coef[j] != 0.0) return false; return true; };
            auto xt_res = [&](int j) {
                const int tj = feat_term(j);
                double s = Gt[tj * KS + tgt];
                for (int k2 = 0; k2 < F; ++k2) s = fma(-Gt[tj * KS + feat_term(k2)], s_coef[k2], s);
                return fma(-s_b, Gt[tj * KS + T], s);
            };
            while (iteration < LASSO_MAX_ITER && !all_zero()) {
                ++s_sweeps;
                double max_change = 0.0;
                // order = argsort(cv)[::-1]: descending CV, ties -> higher index first
                int order[TMAX];
                for (int j = 0; j < F; ++j) order[j] = j;
                for (int a = 1; a < F; ++a) {
                    const int oj = order[a];
                    int b2 = a - 1;
                    while (b2 >= 0 && (s_cv[order[b2]] < s_cv[oj] || (s_cv[order[b2]] == s_cv[oj] && order[b2] < oj))) {
                        order[b2 + 1] = order[b2];
                        --b2;
                    }
                    order[b2 + 1] = oj;
                }
                for (int q = 0; q < F; ++q) {
                    const int j = order[q];
                    const double old_coef = s_coef[j];
                    if (old_coef == 0.0) continue;
                    const int tj = feat_term(j);
                    const double norm_sq = Gt[tj * KS + tj];
                    const double y_sq_sum = yy - 2.0 * s_b * sy + Npts * s_b * s_b;
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
                for (int j = 0; j < F; ++j) max_abs = fmax(max_abs, fabs(s_coef[j]));
                if (max_change <= LASSO_TOL * max_abs) {
                    const double y_sq_sum = yy - 2.0 * s_b * sy + Npts * s_b * s_b;
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
    for (int q = tid; q < W * KS; q += SOLVE_THREADS) {
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
        p.n_active[ev] = nact;
        p.all_zero[ev] = (nact == 0);
        p.iters[ev] = s_sweeps;
        // stability = sum(cv[:-1]) / ndim over the columns of the cached weights (fitness.py:160-169)
        double st = 0.0;
        for (int a = 0; a < n - 1; ++a) st += s_cvraw[a];
        p.stability[ev] = (mode == 2) ? nan("") : st / p.g.ndim;
    }
}


// ------------------------------------------------------------------------------------------
// pass B: residual sums per eval: sum|r|, sum|t|, sum r2^2 with r = t - (F w + b1), r2 = t - (F w + b)
// (fitness.py:135-142 and 70-83).  Persistent CTAs walk over tiles of tile_pts consecutive points
// (double-buffered in shared memory with cp.async).  Each CTA first builds the "residual program" of
// every equation of the launch in shared memory (term ops, weights of every eval, masks); then, per
// tile, its warps take equations from the list dynamically (longest first) and evaluate all evals of
// that equation on the tile, lane <-> point, two points per decode.  Tile sums are added to a per-CTA
// shared table in tile order (fixed order), written once per CTA and reduced by residual_final_kernel.
// ------------------------------------------------------------------------------------------
struct ResidParams {
    long long N;
    const EqDesc* eqs;
    const FactorDesc* factors;
    const int* eq_list;      // equations of this launch (most terms first)
    int n_eq;
    int U;
    int n_tiles;
    int tile_pts;            // points per tile (multiple of 1024)
    const double* tok[MAXU];
    const int* slot_eval;    // [n_eq_total][TMAX] eval id of (equation, target term) or -1
    const double* wterm;     // [n_eval][TMAX]
    const double* intercept; // [n_eval]
    const double* intercept_l1;  // [n_eval]
    double* partial;         // [gridDim.x][n_eq][TMAX][3]
    struct ResidProg* progs; // global: [2 n_eq] sorted programs, then [2 n_eq] staging (residual_prog_kernel)
    int* n_live;             // number of programs with evals (they come first after the sort)
    int split;               // 1: the heaviest programs (NU * NV >= 24) become two programs of half the evals each
    int seg_eqs;             // programs per segment (shared-memory resident)
    int batch_tiles;         // tiles a CTA walks per segment before moving to the next segment
    int nbuf;                // 1 or 2 tile buffers
};

// Per-equation residual program in shared memory (16-byte aligned), COMPACTED: only the NU terms
// some eval needs ("used": targets and surviving features) and only the NV evals that exist.
// The residual of eval dd at a point is  r = sum_u m[dd][u] * c[u] - b1[dd]  with m = +1 at the target
// column and -coefficient at the feature columns, so the inner loop has no data-dependent tests.
struct ResidProg {
    int4 op[TMAX];           // fast terms of the used columns (byte offsets of the slots inside a tile buffer)
    double m[TMAX * TMAX];   // [dd][u]
    double sb[TMAX], sdb[TMAX];
    int term[TMAX];          // used column u -> term index (generic path)
    int foff[TMAX + 1];      // factor ranges by TERM index (generic path)
    int dmap[TMAX];          // compact eval dd -> target term d (row of the sum table)
    int tcol[TMAX];          // compact eval dd -> used column of its target
    int NU, NV, fast, n2;    // n2: used columns with >= 2 slot factors (they come first)
    int n3, le, pad1, pad2;    // n3: depth over the used columns: 0 (<= 2 slot factors), 1 (<= 4), 2 (<= 6); le: list position
    int2 op2[TMAX];            // byte offsets of slots {E, F}
};

static_assert(sizeof(ResidProg) % 16 == 0, "ResidProg is copied in 16-byte pieces");

template <int NU>
__device__ __forceinline__ void residual_point_block(const double (&c)[NU], int NV, unsigned m_a, unsigned sb_a,
                                                     unsigned sdb_a, double (&a1)[NU], double (&a3)[NU], double (&sabs)[NU]) {
#pragma unroll
    for (int u = 0; u < NU; ++u) sabs[u] += fabs(c[u]);
#pragma unroll
    for (int dd = 0; dd < NU; ++dd) {
        if (dd < NV) {
            double r = -lds_f64(sb_a + dd * 8);
#pragma unroll
            for (int u = 0; u < NU; ++u) r = fma(lds_f64(m_a + (dd * TMAX + u) * 8), c[u], r);
            const double q = r + lds_f64(sdb_a + dd * 8);     // L2Fitness keeps the LASSO intercept (fitness.py:80)
            a1[dd] += fabs(r);
            a3[dd] = fma(q, q, a3[dd]);
        }
    }
}

// R points per decode: same arithmetic, weights fetched once for the R points
template <int NU, int R, bool L2>
__device__ __forceinline__ void residual_points(const double (&c)[R][NU], int NV, unsigned m_a, unsigned sb_a,
                                                unsigned sdb_a, double (&a1)[NU], double (&a3)[NU], double (&sabs)[NU]) {
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        if (R == 4) {
            sabs[u] += (fabs(c[0][u]) + fabs(c[1][u])) + (fabs(c[2][u]) + fabs(c[3][u]));
        } else {
            double t = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) t += fabs(c[r][u]);
            sabs[u] += t;
        }
    }
#pragma unroll
    for (int dd = 0; dd < NU; ++dd) {
        if (dd < NV) {
            const double nb = -lds_f64(sb_a + dd * 8);
            double rr[R];
#pragma unroll
            for (int r = 0; r < R; ++r) rr[r] = nb;
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                const double w = lds_f64(m_a + (dd * TMAX + u) * 8);
#pragma unroll
                for (int r = 0; r < R; ++r) rr[r] = fma(w, c[r][u], rr[r]);
            }
            if (!L2) {      // L1 norms only (L2LRFitness, fitness.py:142): no sum of squares
                if (R == 4) a1[dd] += (fabs(rr[0]) + fabs(rr[1])) + (fabs(rr[2]) + fabs(rr[3]));
                else if (R == 2) a1[dd] += fabs(rr[0]) + fabs(rr[1]);
                else a1[dd] += fabs(rr[0]);
                continue;
            }
            const double db = lds_f64(sdb_a + dd * 8);        // L2Fitness keeps the LASSO intercept (fitness.py:80)
            // pairwise (fixed order) instead of a serial chain: the R points are independent
            double q[R];
#pragma unroll
            for (int r = 0; r < R; ++r) q[r] = rr[r] + db;
            if (R == 4) {
                a1[dd] += (fabs(rr[0]) + fabs(rr[1])) + (fabs(rr[2]) + fabs(rr[3]));
                a3[dd] += fma(q[0], q[0], q[1] * q[1]) + fma(q[2], q[2], q[3] * q[3]);
            } else if (R == 2) {
                a1[dd] += fabs(rr[0]) + fabs(rr[1]);
                a3[dd] += fma(q[0], q[0], q[1] * q[1]);
            } else {
                double s1 = 0.0, s3 = a3[dd];
#pragma unroll
                for (int r = 0; r < R; ++r) { s1 += fabs(rr[r]); s3 = fma(q[r], q[r], s3); }
                a1[dd] += s1;
                a3[dd] = s3;
            }
        }
    }
}

template <int NU, bool L2>
__device__ __forceinline__ void residual_reduce(const double (&a1)[NU], const double (&a3)[NU], const double (&sabs)[NU],
                                                const ResidProg* rp, int NV, int lane, double* table_row) {
    double t_abs[NU];
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        double v = sabs[u];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        t_abs[u] = v;
    }
#pragma unroll
    for (int dd = 0; dd < NU; ++dd) {
        if (dd < NV) {
            double v1 = a1[dd], v3 = a3[dd];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                v1 += __shfl_xor_sync(0xffffffffu, v1, off);
                if (L2) v3 += __shfl_xor_sync(0xffffffffu, v3, off);
            }
            if (lane == 0) {
                const int d = rp->dmap[dd], tc = rp->tcol[dd];
                double ta = 0.0;
#pragma unroll
                for (int u = 0; u < NU; ++u) if (u == tc) ta = t_abs[u];
                table_row[d * 3 + 0] += v1;
                table_row[d * 3 + 1] += ta;
                table_row[d * 3 + 2] += v3;
            }
        }
    }
}

// one (tile, equation) task, fast terms; returns through the shared table row
template <int NU, int N2, int DEEP, bool L2>
__device__ __noinline__ void residual_task_fast(unsigned tile, unsigned prog, unsigned ones, const ResidProg* rp, int NV,
                                                int npts, int lane, double* table_row) {
    constexpr int R = (NU <= 6) ? 4 : 2;      // (4 at NU 7, 8 in the L1-only build spills: measured 116 vs 99 ms)
    const unsigned m_a = prog + (unsigned)offsetof(ResidProg, m);
    const unsigned sb_a = prog + (unsigned)offsetof(ResidProg, sb);
    const unsigned sdb_a = prog + (unsigned)offsetof(ResidProg, sdb);
    double a1[NU], a3[NU], sabs[NU];
#pragma unroll
    for (int k = 0; k < NU; ++k) { a1[k] = 0.0; a3[k] = 0.0; sabs[k] = 0.0; }
    TermOffsets<NU, N2, DEEP> off;
    load_term_offsets<NU, N2, DEEP>(prog, prog + (unsigned)offsetof(ResidProg, op2), off);
    const int nfull = npts - npts % (32 * R);
    int i = lane;
    for (; i < nfull; i += 32 * R) {
        double c[R][NU];
        fast_terms_r<NU, N2, DEEP, R, 256>(off, tile + i * 8, ones, c);
        residual_points<NU, R, L2>(c, NV, m_a, sb_a, sdb_a, a1, a3, sabs);
    }
    for (; i < npts; i += 32) {
        double c[1][NU];
        fast_terms_r<NU, N2, DEEP, 1, 256>(off, tile + i * 8, ones, c);
        residual_points<NU, 1, L2>(c, NV, m_a, sb_a, sdb_a, a1, a3, sabs);
    }
    residual_reduce<NU, L2>(a1, a3, sabs, rp, NV, lane, table_row);
}

// same, generic terms (any kind / power / factor count)
template <int NU, bool L2>
__device__ __noinline__ void residual_task_generic(const double* tile_g, int tile_pts, unsigned prog, const ResidProg* rp,
                                                   const FactorDesc* fac, int NV, int npts, int lane, double* table_row) {
    const unsigned m_a = prog + (unsigned)offsetof(ResidProg, m);
    const unsigned sb_a = prog + (unsigned)offsetof(ResidProg, sb);
    const unsigned sdb_a = prog + (unsigned)offsetof(ResidProg, sdb);
    double a1[NU], a3[NU], sabs[NU];
#pragma unroll
    for (int k = 0; k < NU; ++k) { a1[k] = 0.0; a3[k] = 0.0; sabs[k] = 0.0; }
    for (int i = lane; i < npts; i += 32) {
        double c[1][NU];
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            const int t = rp->term[u];
            c[0][u] = eval_term_generic(fac, rp->foff[t], rp->foff[t + 1], tile_g + i, tile_pts);
        }
        residual_points<NU, 1, L2>(c, NV, m_a, sb_a, sdb_a, a1, a3, sabs);
    }
    residual_reduce<NU, L2>(a1, a3, sabs, rp, NV, lane, table_row);
}

template <int NU, bool L2>
__device__ __forceinline__ void residual_task(unsigned tile, const double* tile_g, int tile_pts, unsigned prog, unsigned ones,
                                              const ResidProg* rp, const FactorDesc* fac, int npts, int lane,
                                              double* table_row) {
    if (rp->fast) {
#define RTF(NN, DD) residual_task_fast<NU, (NN), (DD), L2>(tile, prog, ones, rp, rp->NV, npts, lane, table_row)
#if COARSE_CLASSES_B
#define RTD(NN) do { if (rp->n3 == 0) RTF(NN, 0); else RTF(NN, 2); } while (0)
#else
#define RTD(NN) do { if (rp->n3 == 0) RTF(NN, 0); else if (rp->n3 == 1) RTF(NN, 1); else RTF(NN, 2); } while (0)
#endif
        switch (n2_class(rp->n2, COARSE_CLASSES_B)) {
            case 0: RTF(0, 0); break;
#if !COARSE_CLASSES_B
            case 1: RTD(1); break;
#endif
            case 2: if (NU >= 2) { RTD(NU >= 2 ? 2 : NU); break; }
            default: RTD(NU); break;
        }
#undef RTD
#undef RTF
    } else {
        residual_task_generic<NU, L2>(tile_g, tile_pts, prog, rp, fac, rp->NV, npts, lane, table_row);
    }
}

// program kernel (one CTA): builds the residual program of every equation of the launch from the solve outputs and
// stores them in global memory SORTED by code variant (heaviest first), so that a segment of consecutive programs
// runs (almost) one code variant with tasks of similar length.
__global__ void __launch_bounds__(256) residual_prog_kernel(ResidParams p) {
    extern __shared__ int s_key[];                    // [n_eq]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nwarps = blockDim.x >> 5;
    const int TP = p.tile_pts;
    ResidProg* tmp = p.progs + 2 * p.n_eq;            // unsorted staging area behind the sorted programs
    for (int le = warp; le < p.n_eq; le += nwarps) {
        const int e = p.eq_list[le];
        const EqDesc* eq = p.eqs + e;
        // two slots per equation: with few equations per launch the longest (tile, program) task sets the time of
        // every tile, so a program with many evals is cut into two programs over the SAME columns (each eval's
        // arithmetic is unchanged, only the scheduling differs); the second slot stays empty otherwise
        ResidProg* rp = tmp + 2 * le;
        ResidProg* rp2 = rp + 1;
        const int T = eq->T;
        // lane d < T holds the eval id of target d
        const int my_ev = (lane < T) ? p.slot_eval[e * TMAX + lane] : -1;
        const unsigned valid = __ballot_sync(0xffffffffu, my_ev >= 0);
        double wv[2];
        unsigned nz[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int idx = lane + 32 * h;
            const int d = idx >> 3, j = idx & 7;
            const int evd = __shfl_sync(0xffffffffu, my_ev, d);
            wv[h] = (evd >= 0 && j < T) ? p.wterm[(long long)evd * TMAX + j] : 0.0;
            nz[h] = __ballot_sync(0xffffffffu, wv[h] != 0.0);
        }
        unsigned used = valid;
#pragma unroll
        for (int d = 0; d < 4; ++d) used |= ((nz[0] >> (d * 8)) & 0xffu) | ((nz[1] >> (d * 8)) & 0xffu);
        const int NU = __popc(used), NV = __popc(valid);
        const int h0 = (p.split && NV >= 3 && NU * NV >= 24) ? (NV + 1) / 2 : NV;      // evals of the first program
        // used columns in evaluation order (most slot factors first): rank of this lane's term
        int my_u = 0;
        {
            const int my_pos = (lane < T) ? eq->pos[lane] : 0;
            for (int t2 = 0; t2 < T; ++t2)
                if ((used >> t2 & 1u) && eq->pos[t2] < my_pos) ++my_u;
        }
        const int4 my_op = (lane < T) ? eq->fastop[eq->pos[lane]] : make_int4(0, -1, -1, -1);
        const int2 my_op2 = (lane < T) ? eq->fastop2[eq->pos[lane]] : make_int2(-1, -1);
        const bool mine = lane < T && (used >> lane & 1u);
        const unsigned two = __ballot_sync(0xffffffffu, mine && my_op.z >= 0);
        const unsigned three = __ballot_sync(0xffffffffu, mine && my_op.y >= 0);
        const unsigned five = __ballot_sync(0xffffffffu, mine && my_op2.x >= 0);
        for (int h = 0; h < 2; ++h) {
            ResidProg* r = h ? rp2 : rp;
            for (int q = lane; q < TMAX * TMAX; q += 32) r->m[q] = 0.0;
            if (lane < TMAX) {
                r->op[lane] = make_int4(0, -1, -1, -1); r->op2[lane] = make_int2(-1, -1);
                r->sb[lane] = 0.0; r->sdb[lane] = 0.0; r->term[lane] = 0; r->dmap[lane] = 0; r->tcol[lane] = 0;
            }
        }
        __syncwarp();
        if (mine) {
            int4 d = my_op;
            d.x *= TP * 8;
            if (d.y >= 0) d.y *= TP * 8;
            if (d.z >= 0) d.z *= TP * 8;
            if (d.w >= 0) d.w *= TP * 8;
            rp->op[my_u] = d; rp2->op[my_u] = d;
            int2 d2 = my_op2;
            if (d2.x >= 0) d2.x *= TP * 8;
            if (d2.y >= 0) d2.y *= TP * 8;
            rp->op2[my_u] = d2; rp2->op2[my_u] = d2;
            rp->term[my_u] = lane; rp2->term[my_u] = lane;
        }
        if (lane <= TMAX) { rp->foff[lane] = eq->fac_off[lane]; rp2->foff[lane] = eq->fac_off[lane]; }
        if (my_ev >= 0) {
            const int dd_all = __popc(valid & ((1u << lane) - 1u));
            ResidProg* r = dd_all >= h0 ? rp2 : rp;
            const int dd = dd_all >= h0 ? dd_all - h0 : dd_all;
            const double b1 = p.intercept_l1[my_ev];
            r->sb[dd] = b1;
            r->sdb[dd] = b1 - p.intercept[my_ev];
            r->dmap[dd] = lane;
            r->tcol[dd] = my_u;
            r->m[dd * TMAX + my_u] = 1.0;
        }
        __syncwarp();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int idx = lane + 32 * h;
            const int d = idx >> 3, j = idx & 7;
            const int uj = __shfl_sync(0xffffffffu, my_u, j);
            if (wv[h] != 0.0) {
                const int dd_all = __popc(valid & ((1u << d) - 1u));
                if (dd_all >= h0) rp2->m[(dd_all - h0) * TMAX + uj] = -wv[h];
                else rp->m[dd_all * TMAX + uj] = -wv[h];
            }
        }
        if (lane == 0) {
            const int n2 = __popc(two), n3 = five ? 2 : (three ? 1 : 0);
            for (int h = 0; h < 2; ++h) {
                ResidProg* r = h ? rp2 : rp;
                const int nv = h ? NV - h0 : h0;
                r->NU = NU; r->NV = nv; r->fast = eq->fast; r->n2 = n2; r->n3 = n3; r->le = le;
                s_key[2 * le + h] = (nv == 0) ? -1
                                              : ((NU << 12) | (n2_class(n2, COARSE_CLASSES_B) << 9) |
                                                 ((COARSE_CLASSES_B ? (n3 > 0 ? 2 : 0) : n3) << 7) | ((eq->fast ? 0 : 1) << 6) | (nv << 2));
            }
        }
    }
    __syncthreads();
    __threadfence_block();
    if (tid == 0) {
        int live = 0;
        for (int j = 0; j < 2 * p.n_eq; ++j) live += s_key[j] >= 0 ? 1 : 0;
        *p.n_live = live;
    }
    // rank sort (keys descending, ties by slot), then copy every program to its sorted place
    for (int le = warp; le < 2 * p.n_eq; le += nwarps) {
        const int k = s_key[le];
        int cnt = 0;
        for (int j = lane; j < 2 * p.n_eq; j += 32) {
            const int kj = s_key[j];
            cnt += (kj > k || (kj == k && j < le)) ? 1 : 0;
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
        const int4* src = (const int4*)(tmp + le);
        int4* dst = (int4*)(p.progs + cnt);
        for (int q = lane; q < (int)(sizeof(ResidProg) / 16); q += 32) dst[q] = src[q];
    }
}

// main kernel.  Persistent CTAs (one per SM) own a fixed set of tiles (tile b, b + grid, ...), taken in batches of
// `batch_tiles`.  For every batch the CTA walks the sorted program list segment by segment (`seg_eqs` programs,
// copied from global memory into shared memory), and for every segment over the tiles of the batch (double
// buffered cp.async; the re-reads of a batch come from L2): all warps of the CTA therefore execute one code
// variant on tasks of similar length for batch_tiles tiles in a row -- measured 1.5x faster per task than mixing
// all variants on every tile (instruction cache, tail of the slowest warp).  Per equation the tiles are still
// accumulated in increasing order into the CTA's shared table, so the sums do not depend on the segmenting.
template <int NW, bool L2>
__global__ void __launch_bounds__(NW * 32, 1) residual_tile_kernel(ResidParams p) {
    extern __shared__ __align__(16) double smem[];
    __shared__ const double* s_tok[MAXU];
    __shared__ int s_next;
    __shared__ __align__(16) double s_ones[128];     // 4 rows x 256 bytes of 1.0 (fast_terms_r)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int q = tid; q < 128; q += blockDim.x) s_ones[q] = 1.0;
    const int nwarps = blockDim.x >> 5;
    const int TP = p.tile_pts;
    const size_t buf_doubles = (size_t)p.U * TP;
    double* buf0 = smem;                                                       // [nbuf][U][tile_pts]
    ResidProg* progs = (ResidProg*)(smem + (size_t)p.nbuf * buf_doubles);      // [seg_eqs]
    double* table = (double*)(progs + p.seg_eqs);                              // [n_eq][TMAX][3]
    if (tid < MAXU) s_tok[tid] = p.tok[tid];
    for (int q = tid; q < p.n_eq * TMAX * 3; q += blockDim.x) table[q] = 0.0;
    __syncthreads();

    auto issue = [&](int tI, double* dst_buf) {
        const long long start = (long long)tI * TP;
        const int npts = (int)((p.N - start) < TP ? (p.N - start) : TP);
        const int half = TP / 2;
        const int total = p.U * half;
        for (int q = tid; q < total; q += blockDim.x) {
            const int us = q / half, part = q - us * half;
            const int c0 = part * 2;
            const double* src = s_tok[us] + start + c0;
            double* dst = dst_buf + (size_t)us * TP + c0;
            if (c0 + 2 <= npts) cp_async16(dst, src);
            else if (c0 < npts) cp_async8(dst, src);
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    const int n_prog = *p.n_live;
    const int n_seg = (n_prog + p.seg_eqs - 1) / p.seg_eqs;
    const int my_tiles = (p.n_tiles > (int)blockIdx.x) ? (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    int k_buf = 0;                      // buffer that receives / holds the current tile
    for (int t0 = 0; t0 < my_tiles; t0 += p.batch_tiles) {
        const int nb = (my_tiles - t0) < p.batch_tiles ? (my_tiles - t0) : p.batch_tiles;
        for (int sg = 0; sg < n_seg; ++sg) {
            const int e0 = sg * p.seg_eqs;
            const int ne = (n_prog - e0) < p.seg_eqs ? (n_prog - e0) : p.seg_eqs;
            // ---- the segment's programs: global (sorted) -> shared; with one buffer the first tile goes along
            {
                const int4* src = (const int4*)(p.progs + e0);
                int4* dst = (int4*)progs;
                const int n16 = ne * (int)(sizeof(ResidProg) / 16);
                for (int q = tid; q < n16; q += blockDim.x) cp_async16(dst + q, src + q);
                // first tile of the (batch, segment) walk; with two buffers the previous walk may have prefetched it
                if (p.nbuf == 1 || (t0 == 0 && sg == 0)) issue((int)blockIdx.x + t0 * (int)gridDim.x, buf0 + (size_t)k_buf * buf_doubles);
                else asm volatile("cp.async.commit_group;\n" ::: "memory");
            }
            for (int k = 0; k < nb; ++k) {
                double* cur = buf0 + (size_t)k_buf * buf_doubles;
                const int tI = (int)blockIdx.x + (t0 + k) * (int)gridDim.x;
                // ---- prefetch the next tile of the walk into the other buffer
                int nxt = -1;
                if (k + 1 < nb) nxt = tI + (int)gridDim.x;
                else if (sg + 1 < n_seg) nxt = (int)blockIdx.x + t0 * (int)gridDim.x;
                else if (t0 + nb < my_tiles) nxt = (int)blockIdx.x + (t0 + nb) * (int)gridDim.x;
                if (p.nbuf == 2 && nxt >= 0) {
                    issue(nxt, buf0 + (size_t)(k_buf ^ 1) * buf_doubles);
                    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
                } else {
                    if (p.nbuf == 1 && k > 0) issue(tI, cur);
                    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
                }
                if (tid == 0) s_next = nwarps;
                __syncthreads();
                const long long start = (long long)tI * TP;
                const int npts = (int)((p.N - start) < TP ? (p.N - start) : TP);
                const unsigned cur_s = smem_addr(cur);
                int ti = warp;
                while (ti < ne) {
                    const ResidProg* rp = progs + ti;
                    if (rp->NV > 0) {
                        const unsigned prog_s = smem_addr(rp);
                        double* row = table + (size_t)rp->le * TMAX * 3;
#define RT(UU) residual_task<UU, L2>(cur_s, cur, TP, prog_s, smem_addr(s_ones), rp, p.factors, npts, lane, row)
                        switch (rp->NU) {
                            case 1: RT(1); break;
                            case 2: RT(2); break;
                            case 3: RT(3); break;
                            case 4: RT(4); break;
                            case 5: RT(5); break;
                            case 6: RT(6); break;
                            case 7: RT(7); break;
                            default: RT(8); break;
                        }
#undef RT
                    }
                    if (lane == 0) ti = atomicAdd(&s_next, 1);
                    ti = __shfl_sync(0xffffffffu, ti, 0);
                }
                __syncthreads();            // everyone is done with `cur` (and, at k = nb - 1, with the programs)
                if (p.nbuf == 2) k_buf ^= 1;
            }
        }
    }
    double* out = p.partial + (size_t)blockIdx.x * p.n_eq * TMAX * 3;
    for (int q = tid; q < p.n_eq * TMAX * 3; q += blockDim.x) out[q] = table[q];
}

// fixed-order reduction of the per-CTA partial sums -> sums[eval][3]
struct ResidFinalParams {
    const int* eq_list;
    const int* slot_eval;
    const double* partial;
    double* sums;
    int n_eq;
    int n_cta;
};

__global__ void residual_final_kernel(ResidFinalParams p) {
    const int le = blockIdx.x;
    const int e = p.eq_list[le];
    const int tid = threadIdx.x;          // TMAX*3 threads
    if (tid >= TMAX * 3) return;
    const int d = tid / 3;
    const int ev = p.slot_eval[e * TMAX + d];
    if (ev < 0) return;
    double s = 0.0;
    for (int b = 0; b < p.n_cta; ++b) s += p.partial[((size_t)b * p.n_eq + le) * TMAX * 3 + tid];
    p.sums[(long long)ev * 3 + (tid - d * 3)] = s;
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
    AxisPlan axes[EPDE_MAX_NDIM];
    bool geom_ready = false;
    Geom geom;
    int n_cells = 0;             // nA * nB * npieces
    DevBuf d_geom;               // packed geometry tables
    DevBuf d_marg, d_cells, d_seg, d_win, d_enc, d_out, d_partial, d_given, d_progs, d_fd;
    void* h_pinned = nullptr; size_t h_pinned_cap = 0;
    epde_batch* live = nullptr;
    long long launches = 0;
    size_t cell_budget_bytes = (size_t)12 << 30;   // scratch cap for the cell Grams of one launch group
    int sm_count = 148;
    int max_smem = 227 * 1024;
    int resid_warps = RES_WARPS;
    int gram_warps = GRAM_WARPS;
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
    const AxisPlan& lastp = c->axes[last];
    const long long nl = c->shape[last];
    // pieces: sub-segments of the last axis no wider than a tile
    std::vector<int> piece_start, piece_seg;
    for (auto& cl : lastp.cells)
        for (int lo = cl.lo; lo < cl.hi; lo += TILE_COLS) { piece_start.push_back(lo); piece_seg.push_back(cl.seg); }
    piece_start.push_back((int)nl);
    const int npieces = (int)piece_seg.size();
    // chunks: runs of pieces with total width <= TILE_COLS
    struct Chunk { int p_lo, np, col_lo, ncols; };
    std::vector<Chunk> chunks;
    for (int pI = 0; pI < npieces;) {
        int q = pI;
        while (q < npieces && piece_start[q + 1] - piece_start[pI] <= TILE_COLS) ++q;
        chunks.push_back({pI, q - pI, piece_start[pI], piece_start[q] - piece_start[pI]});
        pI = q;
    }
    const int nA = (int)cellsA.size(), nB = (int)cellsB.size();
    std::vector<UnitDesc> units;
    units.reserve((size_t)nA * nB * chunks.size());
    int max_rows = 1;
    for (int a = 0; a < nA; ++a)
        for (int b = 0; b < nB; ++b) {
            max_rows = std::max(max_rows, (cellsA[a].hi - cellsA[a].lo) * (cellsB[b].hi - cellsB[b].lo));
            for (auto& ch : chunks) {
                UnitDesc u;
                memset(&u, 0, sizeof u);
                u.a_lo = cellsA[a].lo; u.a_hi = cellsA[a].hi;
                u.b_lo = cellsB[b].lo; u.b_hi = cellsB[b].hi;
                u.col_lo = ch.col_lo; u.ncols = ch.ncols;
                u.cell = a * nB + b;
                u.piece_lo = ch.p_lo; u.npieces = ch.np;
                u.wlog = -1;
                for (int wl = 0; wl <= 5 && u.wlog < 0; ++wl) {
                    bool ok = true;
                    for (int q = 0; q < ch.np && ok; ++q)
                        ok = (piece_start[ch.p_lo + q + 1] - piece_start[ch.p_lo + q]) == (1 << wl);
                    if (ok) u.wlog = wl;
                }
                units.push_back(u);
            }
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
    g.aligned16 = (nd == 1 || (nl % 2 == 0)) ? 1 : 0;
    g.N = c->N;
    g.units = (const UnitDesc*)(base + o_units);
    g.piece_start = (const int*)(base + o_piece);
    g.marg = (const MargDesc*)(base + o_marg);
    g.win_lo_row = (const int*)(base + o_wlo);
    g.win_hi_row = (const int*)(base + o_whi);
    c->n_cells = nA * nB * npieces;
    c->geom_ready = true;
    return 0;
}

struct LaunchGroup {             // equations evaluated together: <= MAXU distinct tokens, cell table within budget
    std::vector<int> eqs;        // most terms first
    std::vector<int> token_ids;  // slot -> token id
    int list_off = 0;            // offset of the group's equation list inside d_lists
    long long cell_doubles = 0;
};

struct epde_batch {
    epde_ctx* ctx = nullptr;
    int n_eq = 0, n_eval = 0, n_fac = 0;
    std::vector<EqDesc> eqs;
    std::vector<LaunchGroup> groups;
    std::vector<int> evals_gs8, evals_gs16;
    // device pointers into ctx->d_enc
    EqDesc* d_eqs = nullptr;
    FactorDesc* d_factors = nullptr;
    int* d_eval_eq = nullptr; int* d_eval_tgt = nullptr; int* d_slot_eval = nullptr;
    int* d_lists = nullptr;                      // equation lists by group, eval lists by GS
    int off_gs8 = 0, off_gs16 = 0;
    // outputs in ctx->d_out
    double* d_coef = nullptr; double* d_intercept = nullptr; double* d_stab = nullptr;
    double* d_sw = nullptr; double* d_wterm = nullptr; double* d_sums = nullptr; double* d_b1 = nullptr;
    int* d_nact = nullptr; int* d_allzero = nullptr; int* d_iters = nullptr;
    size_t out_bytes = 0;
    long long seg_doubles = 0, win_doubles = 0;
    int* d_mode = nullptr; double* d_given_w = nullptr; double* d_given_b = nullptr;
    bool solved = false;
    bool want_l2 = true;       // epde_batch_set_norms
};

static size_t gram_smem_bytes(const epde_ctx* c, int U, int nwarps, int n_eq) {
    (void)nwarps;
    return sizeof(double) * ((size_t)U * c->geom.max_rows * TILE_COLS) + (size_t)n_eq * sizeof(WarpProg);
}

static int launch_gram_group(epde_ctx* c, epde_batch* b, const LaunchGroup& g, double* cells) {
    GramParams p;
    memset(&p, 0, sizeof p);
    p.g = c->geom; p.eqs = b->d_eqs; p.factors = b->d_factors;
    p.eq_list = b->d_lists + g.list_off; p.n_eq = (int)g.eqs.size();
    p.U = (int)g.token_ids.size();
    for (int u = 0; u < p.U; ++u) p.tok[u] = c->tokens[g.token_ids[u]];
    p.cells = cells;
    // persistent CTAs (one per SM) walk over the units; the programs of the launch's equations live in shared
    // memory next to the tile, so a group whose programs do not fit is split into several launches
    const int ne_all = (int)g.eqs.size();
    const size_t fixed = gram_smem_bytes(c, p.U, c->gram_warps, 0);
    if (fixed + sizeof(WarpProg) > (size_t)c->max_smem)
        return fail("pass A tile needs %zu bytes of shared memory (> %d)", fixed + sizeof(WarpProg), c->max_smem);
    const int cap = (int)(((size_t)c->max_smem - fixed) / sizeof(WarpProg));
    const int n_launch = (ne_all + cap - 1) / cap;
    const int per_launch = (ne_all + n_launch - 1) / n_launch;
    for (int e0 = 0; e0 < ne_all; e0 += per_launch) {
        p.eq_list = b->d_lists + g.list_off + e0;
        p.n_eq = std::min(per_launch, ne_all - e0);
        const int nwarps = std::max(1, std::min(c->gram_warps, p.n_eq));
        const size_t smem = gram_smem_bytes(c, p.U, nwarps, p.n_eq);
        const unsigned grid = (unsigned)std::min(c->geom.n_units, c->sm_count);
        {
            KernelSpan ks(c, EPDE_KERNEL_GRAM);
            gram_tile_kernel<GRAM_WARPS><<<grid, nwarps * 32, smem, c->stream>>>(p);
        }
        CK(cudaGetLastError());
    }
    return 0;
}

static int launch_resid_group(epde_ctx* c, epde_batch* b, const LaunchGroup& g, double* sums) {
    const int U = (int)g.token_ids.size();
    // tile = 1024 points (measured at 512^3, population 128: 2048-point tiles are slower); two buffers when they
    // fit next to one segment of programs and the sum table
    int tile_pts = RES_TILE_MIN;
    if (const char* ev = getenv("EPDE_RES_TILE")) { const int v = atoi(ev); if (v == 1024 || v == 2048 || v == 4096) tile_pts = v; }
    int seg_eqs = 64, batch_tiles = 4;      // measured at 512^3 / 128 equations: (64, 4) 117 ms, (16, 8) 129, (8, 8) 143
    if (const char* ev = getenv("EPDE_RES_SEG")) seg_eqs = std::max(1, atoi(ev));
    if (const char* ev = getenv("EPDE_RES_BATCH")) batch_tiles = std::max(1, atoi(ev));
    const long long n_tiles = (c->N + tile_pts - 1) / tile_pts;
    const int ne_all = (int)g.eqs.size();
    const size_t tile_bytes = sizeof(double) * (size_t)U * tile_pts;
    const size_t table_eq = sizeof(double) * TMAX * 3;
    // equations per launch: the sum table of all of them stays in shared memory
    const size_t fixed1 = tile_bytes + (size_t)seg_eqs * sizeof(ResidProg);
    if (fixed1 + table_eq > (size_t)c->max_smem) return fail("pass B tiles need %zu bytes of shared memory", fixed1 + table_eq);
    const int cap = (int)(((size_t)c->max_smem - fixed1) / table_eq);
    const int n_launch = (ne_all + cap - 1) / cap;
    const int per_launch = (ne_all + n_launch - 1) / n_launch;
    for (int e0 = 0; e0 < ne_all; e0 += per_launch) {
        const int ne = std::min(per_launch, ne_all - e0);
        // few programs per CTA (<= 5 per warp): halve the longest tasks (residual_prog_kernel).  Measured: 32 equations
        // per launch, slowest of 8 population shards 8.0 -> 5.5 ms; at 64 per launch the recomputed columns cost more
        // than the tail (512^3: 91 -> 103 ms), hence the limit
        const int split = ne <= 40 ? 1 : 0;
        const int nwarps = std::max(1, std::min(c->resid_warps, std::min(split ? 2 * ne : ne, seg_eqs)));
        ResidParams p;
        memset(&p, 0, sizeof p);
        p.N = c->N; p.eqs = b->d_eqs; p.factors = b->d_factors;
        p.eq_list = b->d_lists + g.list_off + e0; p.n_eq = ne; p.U = U; p.n_tiles = (int)n_tiles; p.tile_pts = tile_pts;
        for (int u = 0; u < U; ++u) p.tok[u] = c->tokens[g.token_ids[u]];
        p.slot_eval = b->d_slot_eval; p.wterm = b->d_wterm; p.intercept = b->d_intercept; p.intercept_l1 = b->d_b1;
        p.seg_eqs = std::min(seg_eqs, split ? 2 * ne : ne); p.batch_tiles = batch_tiles; p.split = split;
        const size_t rest = (size_t)p.seg_eqs * sizeof(ResidProg) + (size_t)ne * table_eq;
        p.nbuf = (2 * tile_bytes + rest <= (size_t)c->max_smem) ? 2 : 1;
        const size_t smem = p.nbuf * tile_bytes + rest;
        if (c->d_progs.ensure(4 * (size_t)ne * sizeof(ResidProg) + 16)) return 1;
        p.progs = (ResidProg*)c->d_progs.ptr;
        p.n_live = (int*)(p.progs + 4 * (size_t)ne);
        // one persistent CTA per SM: the tile -> CTA assignment (and with it the summation order) depends only on
        // the grid size, never on what else is in the batch
        const int grid = (int)std::min<long long>(n_tiles, (long long)c->sm_count);
        if (c->d_partial.ensure(sizeof(double) * (size_t)grid * ne * TMAX * 3)) return 1;
        p.partial = (double*)c->d_partial.ptr;
        {
            KernelSpan ks(c, EPDE_KERNEL_RESIDUAL);
            residual_prog_kernel<<<1, 256, sizeof(int) * 2 * (size_t)ne, c->stream>>>(p);
        }
        CK(cudaGetLastError());
        {
            KernelSpan ks(c, EPDE_KERNEL_RESIDUAL);
            if (b->want_l2) residual_tile_kernel<RES_WARPS, true><<<grid, nwarps * 32, smem, c->stream>>>(p);
            else residual_tile_kernel<RES_WARPS, false><<<grid, nwarps * 32, smem, c->stream>>>(p);
        }
        CK(cudaGetLastError());
        ResidFinalParams fp;
        fp.eq_list = p.eq_list; fp.slot_eval = b->d_slot_eval; fp.partial = p.partial; fp.sums = sums;
        fp.n_eq = ne; fp.n_cta = grid;
        { KernelSpan ks(c, EPDE_KERNEL_RESIDUAL); residual_final_kernel<<<ne, 32, 0, c->stream>>>(fp); }
        CK(cudaGetLastError());
    }
    return 0;
}
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

__global__ void __launch_bounds__(256) fd_gradient_kernel(FdParams p) {
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < p.total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int i = (int)((idx / p.stride) % p.n);
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

__global__ void __launch_bounds__(256) crop_interior_kernel(CropParams p) {
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < p.n_inner;
         idx += (long long)gridDim.x * blockDim.x) {
        long long rem = idx, src = 0;
        for (int d = p.ndim - 1; d >= 0; --d) {
            const long long q = rem % p.inner[d];
            rem /= p.inner[d];
            src += (q + p.off[d]) * p.full_stride[d];
        }
        p.out[idx] = p.in[src];
    }
}

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
extern "C" {

const char* epde_last_error(void) { return g_err.c_str(); }
const char* epde_version(void) { return "epde_b200 0.2.0 sm_100a"; }

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
        // dynamic shared memory limit = opt-in maximum minus the kernels' static shared memory
        const void* fns[3] = {(const void*)gram_tile_kernel<GRAM_WARPS>, (const void*)residual_tile_kernel<RES_WARPS, true>,
                              (const void*)residual_tile_kernel<RES_WARPS, false>};
        int dyn_max = c->max_smem;
        for (int k = 0; k < 3; ++k) {
            cudaFuncAttributes fa;
            cudaError_t fe = cudaFuncGetAttributes(&fa, fns[k]);
            if (fe == cudaSuccess) {
                const int dyn = c->max_smem - (int)fa.sharedSizeBytes;
                fe = cudaFuncSetAttribute(fns[k], cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
                dyn_max = std::min(dyn_max, dyn);
            }
            if (fe != cudaSuccess) {
                delete c;
                return fail("epde_ctx_create: cannot configure shared memory (%s); is this an sm_100 device?",
                            cudaGetErrorString(fe));
            }
        }
        c->max_smem = dyn_max;
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
    c->d_geom.release(); c->d_marg.release(); c->d_cells.release(); c->d_seg.release(); c->d_win.release(); c->d_enc.release();
    c->d_out.release(); c->d_partial.release(); c->d_given.release(); c->d_progs.release(); c->d_fd.release();
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

int epde_cache_upload(epde_ctx* c, int id, const double* host, int64_t n) {
    if (!c || !host) return fail("epde_cache_upload: null argument");
    if (n != c->N) return fail("epde_cache_upload: token %d has %lld values, context expects %lld", id, (long long)n, c->N);
    if (id < 0 || id >= EPDE_MAX_TOKENS) return fail("epde_cache_upload: token id %d out of range", id);
    CK(cudaSetDevice(c->device));
    double*& p = c->tokens[id];
    if (!p) {
        cudaError_t e = cudaMalloc((void**)&p, sizeof(double) * (size_t)n);
        if (e != cudaSuccess) { c->tokens.erase(id); return fail("epde_cache_upload: out of device memory (%s)", cudaGetErrorString(e)); }
    }
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
    if (!p) {
        cudaError_t e = cudaMalloc((void**)&p, sizeof(double) * (size_t)n);
        if (e != cudaSuccess) { c->tokens.erase(id); return fail("epde_cache_upload_device: out of device memory (%s)", cudaGetErrorString(e)); }
    }
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
    for (int64_t i = 0; i < n; ++i)
        if (host[i] != 1.0)
            return fail("epde_ctx_set_sample_weights: only the default all-ones g_func mask is supported "
                        "(value %g at %lld); non-trivial sample weights are not implemented", host[i], (long long)i);
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
    const int n_terms = term_off[n_eq];
    b->n_fac = fac_off[n_terms];
    std::vector<FactorDesc> factors(b->n_fac);
    b->eqs.resize(n_eq);
    const int W = c->geom.n_windows;
    long long s_off = 0, w_off = 0;
    for (int e = 0; e < n_eq; ++e) {
        const int T = term_off[e + 1] - term_off[e];
        if (T < 1 || T > TMAX) { delete b; return fail("equation %d has %d terms; supported: 1..%d", e, T, TMAX); }
        EqDesc& q = b->eqs[e];
        memset(&q, 0, sizeof q);
        q.T = T;
        for (int t = 0; t <= T; ++t) q.fac_off[t] = fac_off[term_off[e] + t];
        for (int t = T + 1; t <= TMAX; ++t) q.fac_off[t] = q.fac_off[T];
        for (int t = 0; t < T; ++t) {
            const int nf = q.fac_off[t + 1] - q.fac_off[t];
            if (nf < 1 || nf > FMAXF) { delete b; return fail("equation %d term %d has %d factors; supported: 1..%d", e, t, nf, FMAXF); }
        }
        q.s_off = s_off;
        s_off += (long long)c->geom.n_rows * (nacc_of(T) + 1);
        q.w_off = w_off;
        w_off += (long long)(W + 1) * (nacc_of(T) + 1);
    }
    b->seg_doubles = s_off;
    b->win_doubles = w_off;
    for (int f = 0; f < b->n_fac; ++f) {
        if (!c->tokens.count(fac_token[f])) { delete b; return fail("factor %d references token %d which is not in the cache", f, fac_token[f]); }
        FactorDesc& d = factors[f];
        memset(&d, 0, sizeof d);
        d.power = fac_power[f];
        d.kind = fac_kind ? fac_kind[f] : 0;
        d.param = fac_param ? fac_param[f] : 0.0;
        if (d.kind < 0 || d.kind > 3) { delete b; return fail("factor %d has unknown kind %d", f, d.kind); }
        d.ipow = (d.power == 1.0) ? 1 : (d.power == 2.0) ? 2 : (d.power == 3.0) ? 3 : 0;
    }
    // ---- launch groups: a group shares one staged tile of <= MAXU distinct tokens.  Equations are taken most
    //      terms first and placed first-fit into the open groups (token count and cell scratch permitting).
    std::vector<int> order(n_eq);
    for (int e = 0; e < n_eq; ++e) order[e] = e;
    std::stable_sort(order.begin(), order.end(), [&](int a, int bb) { return b->eqs[a].T > b->eqs[bb].T; });
    std::vector<std::set<int>> group_tok;
    for (int e : order) {
        const EqDesc& q = b->eqs[e];
        std::set<int> mine;
        for (int f = q.fac_off[0]; f < q.fac_off[q.T]; ++f) mine.insert(fac_token[f]);
        if ((int)mine.size() > MAXU) { delete b; return fail("equation %d references %d distinct tokens; supported: %d", e, (int)mine.size(), MAXU); }
        const long long cd = (long long)c->n_cells * nacc_of(q.T);
        int gi = -1;
        for (size_t k = 0; k < b->groups.size() && gi < 0; ++k) {
            std::set<int> uni = group_tok[k];
            uni.insert(mine.begin(), mine.end());
            if ((int)uni.size() <= MAXU &&
                (size_t)(b->groups[k].cell_doubles + cd) * sizeof(double) <= c->cell_budget_bytes) gi = (int)k;
        }
        if (gi < 0) { b->groups.push_back(LaunchGroup()); group_tok.push_back(std::set<int>()); gi = (int)b->groups.size() - 1; }
        group_tok[gi].insert(mine.begin(), mine.end());
        b->groups[gi].eqs.push_back(e);
        b->groups[gi].cell_doubles += cd;
    }
    // per group: slots, cell offsets, fast-path encoding
    for (size_t gi = 0; gi < b->groups.size(); ++gi) {
        LaunchGroup& g = b->groups[gi];
        g.token_ids.assign(group_tok[gi].begin(), group_tok[gi].end());
        std::map<int, int> slot;
        for (size_t u = 0; u < g.token_ids.size(); ++u) slot[g.token_ids[u]] = (int)u;
        long long c_off = 0;
        for (int e : g.eqs) {
            EqDesc& q = b->eqs[e];
            q.group = (int)gi;
            q.c_off = c_off;
            c_off += (long long)c->n_cells * nacc_of(q.T);
            // fast equation: every term is a product of <= 6 token values once integer powers are expanded
            int nsf[TMAX];
            int sf[TMAX][6];
            q.fast = 1;
            for (int t = 0; t < q.T; ++t) {
                nsf[t] = 0;
                for (int f = q.fac_off[t]; f < q.fac_off[t + 1]; ++f) {
                    factors[f].slot = slot[fac_token[f]];
                    if (factors[f].kind != 0 || factors[f].ipow == 0) { q.fast = 0; continue; }
                    for (int k = 0; k < factors[f].ipow; ++k) {
                        if (nsf[t] < 6) sf[t][nsf[t]] = factors[f].slot;
                        ++nsf[t];
                    }
                }
                if (nsf[t] > 6) q.fast = 0;
            }
            for (int t = 0; t < TMAX; ++t) {
                q.fastop[t] = make_int4(0, -1, -1, -1); q.fastop2[t] = make_int2(-1, -1);
                q.pos[t] = q.perm[t] = (unsigned char)t;
            }
            // evaluation order: most slot factors first (fast equations only)
            q.n2 = q.n3 = 0;
            if (q.fast) {
                int k = 0;
                for (int want = 6; want >= 1; --want)
                    for (int t = 0; t < q.T; ++t)
                        if (nsf[t] == want) { q.perm[k] = (unsigned char)t; q.pos[t] = (unsigned char)k; ++k; }
                for (int t = 0; t < q.T; ++t) {
                    if (nsf[t] >= 2) ++q.n2;
                    if (nsf[t] >= 3 && q.n3 < 1) q.n3 = 1;
                    if (nsf[t] >= 5) q.n3 = 2;
                    int4& op = q.fastop[q.pos[t]];
                    op.x = sf[t][0];
                    if (nsf[t] >= 2) op.z = sf[t][1];
                    if (nsf[t] >= 3) op.y = sf[t][2];
                    if (nsf[t] >= 4) op.w = sf[t][3];
                    if (nsf[t] >= 5) q.fastop2[q.pos[t]].x = sf[t][4];
                    if (nsf[t] >= 6) q.fastop2[q.pos[t]].y = sf[t][5];
                }
            }
            {
                int k = 0;
                for (int i = 0; i < q.T; ++i)
                    for (int j = i; j < q.T; ++j) q.accmap[k++] = (unsigned char)gidx(q.perm[i], q.perm[j], q.T);
                for (int i = 0; i < q.T; ++i) q.accmap[k++] = (unsigned char)(q.T * (q.T + 1) / 2 + q.perm[i]);
            }
        }
    }
    // task order inside a group: most terms first, equations that run the same code variant (T, factor class)
    // next to each other, so the warps of a CTA execute the same instructions at the same time
    for (auto& g : b->groups) {
        auto key = [&](int e) {
            const EqDesc& q = b->eqs[e];
            return (q.T << 8) | ((q.n2 <= 2 ? q.n2 : 3) << 4) | (q.n3 << 2) | (q.fast ? 0 : 2);
        };
        std::stable_sort(g.eqs.begin(), g.eqs.end(), [&](int a, int bb) { return key(a) > key(bb); });
    }
    std::vector<int> slot_eval((size_t)n_eq * TMAX, -1);
    for (int v = 0; v < n_eval; ++v) {
        const int e = eval_eq[v], t = eval_target[v];
        if (e < 0 || e >= n_eq) { delete b; return fail("eval %d references equation %d", v, e); }
        if (t < 0 || t >= b->eqs[e].T) { delete b; return fail("eval %d: target %d outside equation %d (%d terms)", v, t, e, b->eqs[e].T); }
        if (slot_eval[(size_t)e * TMAX + t] >= 0) { delete b; return fail("eval %d duplicates (equation %d, target %d)", v, e, t); }
        slot_eval[(size_t)e * TMAX + t] = v;
        if (b->eqs[e].T + 1 <= 8) b->evals_gs8.push_back(v); else b->evals_gs16.push_back(v);
    }
    // lists: equations by group, evals by solver group size
    std::vector<int> lists;
    for (auto& g : b->groups) {
        g.list_off = (int)lists.size();
        lists.insert(lists.end(), g.eqs.begin(), g.eqs.end());
    }
    b->off_gs8 = (int)lists.size();
    lists.insert(lists.end(), b->evals_gs8.begin(), b->evals_gs8.end());
    b->off_gs16 = (int)lists.size();
    lists.insert(lists.end(), b->evals_gs16.begin(), b->evals_gs16.end());
    if (lists.empty()) lists.push_back(0);
    // pack the encoding into one pinned staging buffer -> one H2D copy
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_eqs = 0;
    const size_t o_fac = align(o_eqs + sizeof(EqDesc) * n_eq);
    const size_t o_eveq = align(o_fac + sizeof(FactorDesc) * std::max(1, b->n_fac));
    const size_t o_evtg = align(o_eveq + sizeof(int) * std::max(1, n_eval));
    const size_t o_slot = align(o_evtg + sizeof(int) * std::max(1, n_eval));
    const size_t o_list = align(o_slot + sizeof(int) * slot_eval.size());
    const size_t total = align(o_list + sizeof(int) * lists.size());
    // the pinned buffer may still feed an earlier asynchronous copy
    cudaStreamSynchronize(c->stream);
    if (ensure_pinned(c, total) || c->d_enc.ensure(total)) { delete b; return 1; }
    char* h = (char*)c->h_pinned;
    memcpy(h + o_eqs, b->eqs.data(), sizeof(EqDesc) * n_eq);
    memcpy(h + o_fac, factors.data(), sizeof(FactorDesc) * b->n_fac);
    if (n_eval) { memcpy(h + o_eveq, eval_eq, sizeof(int) * n_eval); memcpy(h + o_evtg, eval_target, sizeof(int) * n_eval); }
    memcpy(h + o_slot, slot_eval.data(), sizeof(int) * slot_eval.size());
    memcpy(h + o_list, lists.data(), sizeof(int) * lists.size());
    cudaError_t ce = cudaMemcpyAsync(c->d_enc.ptr, h, total, cudaMemcpyHostToDevice, c->stream);
    if (ce != cudaSuccess) { delete b; return fail("encoding upload failed: %s", cudaGetErrorString(ce)); }
    char* d = (char*)c->d_enc.ptr;
    b->d_eqs = (EqDesc*)(d + o_eqs);
    b->d_factors = (FactorDesc*)(d + o_fac);
    b->d_eval_eq = (int*)(d + o_eveq);
    b->d_eval_tgt = (int*)(d + o_evtg);
    b->d_slot_eval = (int*)(d + o_slot);
    b->d_lists = (int*)(d + o_list);
    // outputs
    const size_t ne = (size_t)std::max(1, n_eval);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align(o + bytes); return r; };
    const size_t oc = take(ne * TMAX * 8), oi = take(ne * 8), ost = take(ne * 8), osw = take(ne * W * KS * 8),
                 owt = take(ne * TMAX * 8), osu = take(ne * 3 * 8), ob1 = take(ne * 8), ona = take(ne * 4), oaz = take(ne * 4), oit = take(ne * 4);
    b->out_bytes = o;
    if (c->d_out.ensure(o)) { delete b; return 1; }
    char* od = (char*)c->d_out.ptr;
    b->d_coef = (double*)(od + oc); b->d_intercept = (double*)(od + oi); b->d_stab = (double*)(od + ost);
    b->d_sw = (double*)(od + osw); b->d_wterm = (double*)(od + owt); b->d_sums = (double*)(od + osu); b->d_b1 = (double*)(od + ob1);
    b->d_nact = (int*)(od + ona); b->d_allzero = (int*)(od + oaz); b->d_iters = (int*)(od + oit);
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

int epde_batch_gram(epde_batch* b, double* dev_seg) {
    if (!b || !b->ctx) return fail("null batch");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    if (!dev_seg) {
        if (c->d_seg.ensure(sizeof(double) * (size_t)b->seg_doubles)) return 1;
        dev_seg = (double*)c->d_seg.ptr;
    }
    long long need = 0;
    for (auto& g : b->groups) need = std::max(need, g.cell_doubles);
    if (c->d_cells.ensure(sizeof(double) * (size_t)need)) return 1;
    for (auto& g : b->groups) {
        if (launch_gram_group(c, b, g, (double*)c->d_cells.ptr)) return 1;
        MargParams mp;
        mp.g = c->geom; mp.eqs = b->d_eqs; mp.eq_list = b->d_lists + g.list_off;
        mp.cells = (const double*)c->d_cells.ptr; mp.seg = dev_seg;
        dim3 grid(c->geom.n_rows, (unsigned)g.eqs.size());
        const size_t slab_smem = sizeof(double) * (size_t)c->geom.npieces * nacc_of(TMAX);
        if (c->ndim == 3 && c->geom.nA >= 4 && slab_smem <= 40 * 1024) {
            MargSlabParams sp;
            sp.g = c->geom; sp.eqs = mp.eqs; sp.eq_list = mp.eq_list; sp.cells = mp.cells; sp.seg = dev_seg;
            sp.eq_stride = (long long)c->geom.nA * (1 + c->geom.nB + c->geom.npieces) * nacc_of(TMAX);
            if (c->d_marg.ensure(sizeof(double) * (size_t)sp.eq_stride * g.eqs.size())) return 1;
            sp.part = (double*)c->d_marg.ptr;
            KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
            marginal_slab_kernel<<<dim3(c->geom.nA, (unsigned)g.eqs.size()), 256, slab_smem, c->stream>>>(sp);
            marginal_rows_kernel<<<grid, 64, 0, c->stream>>>(sp);
            c->launches++;
        } else {
            KernelSpan ks(c, EPDE_KERNEL_MARGINAL);
            marginal_kernel<<<grid, 256, 0, c->stream>>>(mp);
        }
        CK(cudaGetLastError());
    }
    return 0;
}

int epde_batch_solve(epde_batch* b, const double* dev_seg) {
    if (!b || !b->ctx) return fail("null batch");
    epde_ctx* c = b->ctx;
    CK(cudaSetDevice(c->device));
    double* seg = dev_seg ? (double*)dev_seg : (double*)c->d_seg.ptr;
    if (!seg) return fail("epde_batch_solve: no segment Grams (call epde_batch_gram first)");
    if (c->d_win.ensure(sizeof(double) * (size_t)b->win_doubles)) return 1;
    WindowParams wp;
    wp.g = c->geom; wp.eqs = b->d_eqs; wp.seg = seg; wp.win = (double*)c->d_win.ptr;
    { KernelSpan ks(c, EPDE_KERNEL_MARGINAL); window_gram_kernel<<<dim3(c->geom.n_windows + 1, b->n_eq), 64, 0, c->stream>>>(wp); }
    CK(cudaGetLastError());
    if (b->n_eval > 0) {
        SolveParams sp;
        sp.g = c->geom; sp.eqs = b->d_eqs; sp.eval_eq = b->d_eval_eq; sp.eval_tgt = b->d_eval_tgt; sp.win = (const double*)c->d_win.ptr;
        sp.mode = b->d_mode; sp.given_wterm = b->d_given_w; sp.given_b = b->d_given_b;
        sp.coef = b->d_coef; sp.intercept = b->d_intercept; sp.intercept_l1 = b->d_b1; sp.stability = b->d_stab; sp.sw = b->d_sw;
        sp.wterm = b->d_wterm; sp.n_active = b->d_nact; sp.all_zero = b->d_allzero; sp.iters = b->d_iters;
        const size_t smem = sizeof(double) * (size_t)c->geom.n_windows * KS;
        if (!b->evals_gs8.empty()) {
            sp.eval_list = b->d_lists + b->off_gs8;
            { KernelSpan ks(c, EPDE_KERNEL_SOLVE); lasso_solve_kernel<8><<<(unsigned)b->evals_gs8.size(), SOLVE_THREADS, smem, c->stream>>>(sp); }
            CK(cudaGetLastError());
        }
        if (!b->evals_gs16.empty()) {
            sp.eval_list = b->d_lists + b->off_gs16;
            { KernelSpan ks(c, EPDE_KERNEL_SOLVE); lasso_solve_kernel<16><<<(unsigned)b->evals_gs16.size(), SOLVE_THREADS, smem, c->stream>>>(sp); }
            CK(cudaGetLastError());
        }
    }
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
    for (auto& g : b->groups)
        if (launch_resid_group(c, b, g, sums)) return 1;
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
    if (ensure_pinned(c, std::max(b->out_bytes, (size_t)1))) return 1;
    CK(cudaMemcpyAsync(c->h_pinned, c->d_out.ptr, b->out_bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    const char* h = (const char*)c->h_pinned;
    const char* d0 = (const char*)c->d_out.ptr;
    auto hp = [&](const void* dptr) { return h + ((const char*)dptr - d0); };
    const int W = c->geom.n_windows;
    if (coef) memcpy(coef, hp(b->d_coef), sizeof(double) * ne * TMAX);
    if (intercept) memcpy(intercept, hp(b->d_intercept), sizeof(double) * ne);
    if (stability) memcpy(stability, hp(b->d_stab), sizeof(double) * ne);
    if (sw_weights) memcpy(sw_weights, hp(b->d_sw), sizeof(double) * (size_t)ne * W * KS);
    if (n_active) memcpy(n_active, hp(b->d_nact), sizeof(int) * ne);
    if (all_zero) memcpy(all_zero, hp(b->d_allzero), sizeof(int) * ne);
    if (iters) memcpy(iters, hp(b->d_iters), sizeof(int) * ne);
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
    fd_gradient_kernel<<<grid, 256, 0, c->stream>>>(p);
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
    if (!tok) {
        cudaError_t e = cudaMalloc((void**)&tok, sizeof(double) * (size_t)c->N);
        if (e != cudaSuccess) { c->tokens.erase(id); return fail("epde_cache_upload_interior: out of device memory (%s)", cudaGetErrorString(e)); }
    }
    p.out = tok;
    const int grid = (int)std::min<long long>((c->N + 255) / 256, (long long)c->sm_count * 16);
    c->launches++;
    crop_interior_kernel<<<grid, 256, 0, c->stream>>>(p);
    CK(cudaGetLastError());
    return 0;
}

}  // extern "C"
