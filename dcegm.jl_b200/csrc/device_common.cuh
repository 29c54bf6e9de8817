// device_common.cuh — shared-memory layout, polyline primitives and block-level helpers used by
// the solver kernels.  FP64 throughout; compiled with -fmad=false so that every +,-,*,/ below is
// a separately rounded IEEE operation in the order written, and pow/exp/log come from dc_math.h
// (IEEE ops + fma only) instead of the CUDA math library.  The CPU oracle, compiled with
// -ffp-contract=off, reproduces every kernel result bit for bit in its `portable` mode: decisions
// such as `x[k+1] > x[k]` or the arg-max over segments see the same bits.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dcegm.h"
#include "host_setup.hpp"
#include "dc_math.h"

namespace dcegm {

#ifdef DC_T512
constexpr int kThreads = 512;
constexpr int kCtasPerSm = 1;
#elif defined(DC_CTAS3)
constexpr int kThreads = 256;
constexpr int kCtasPerSm = 3;
#else
constexpr int kThreads = 256;
constexpr int kCtasPerSm = 2;
#endif   // threads per CTA of the persistent solver kernels
constexpr int kWarps = kThreads / 32;
constexpr int kMaxNy = 64;     // income nodes staged in shared memory by solve_f_kernel
constexpr int kSegCap = 192;    // max sections of one splitLine (shared-memory resident solvers)
constexpr int kSwitchMaxIsect = 8;  // max intersections produced by one change of winner

// geometry of the resident solution; must match oracle_layout() (tests compare slot by slot)
__host__ __device__ inline void make_layout(int model, int nmodels, int na, int ny, int nT, int keep_dirty,
                                            int has_lines, dcegm_layout* l) {
    l->model = model;
    l->nmodels = nmodels;
    l->na = na;
    l->ny = ny;
    l->nT = nT;
    l->nstate = model == DCEGM_MODEL_G ? ny : 1;
    l->isec_cap = na / 16 > 32 ? na / 16 : 32;
    int cap = na + na / 10 + 2 * l->isec_cap + 2;
    l->cap = (cap + 7) / 8 * 8;
    l->rem_words = (l->cap + 31) / 32;
    l->seg_cap = kSegCap;
    l->has_dirty = keep_dirty;
    l->has_lines = has_lines;
}

// device pointers of one batch solution + inputs
struct DeviceBatch {
    dcegm_layout lay;
    const DevModel* models;
    const double* grids;
    const double* off;
    const double* wgt;
    double2* lines;          // [nmodels][nT][nstate][2][2][cap]
    int32_t* lens;           // [nmodels][nT][nstate][2][2]
    double* vbound;          // [nmodels][nT][nstate][2]
    dcegm_env_meta* meta;    // [nmodels][nT][nstate]
    double2* isects;         // [nmodels][nT][nstate][isec_cap]
    uint32_t* removed;       // [nmodels][nT][nstate][rem_words]
    double2* dirty;          // [nmodels][nT][nstate][2][2][na] or null
    int32_t* status;         // [nmodels]
    dcegm_summary* summaries;  // [nmodels]
    int first;               // first model of this launch (grid = number of models of the launch)
    const int32_t* order;    // launch order of a whole-batch solve: CTA b solves model order[b] (null: model first + b)
    // pipelined solve-to-host in ONE launch: a finished model adds 1 to chunk_done[model_chunk[m]] (after a fence), the
    // packing stream waits for a chunk's counter to reach the chunk's size (stream memory operation); null otherwise
    int32_t* chunk_done;
    const uint16_t* model_chunk;
};

// inputs / outputs of the forward simulator (solver_sim.cu)
struct SimBatch {
    const SimModel* sim;
    const double* yvec;  // [nmodels][ny]
    const double *wsh, *dsh, *rsh, *w0sh;  // given shocks (device) or null
    const int32_t* y0;
    int given;           // 1: shocks given, all models share nsims_given
    int nsims_given;
    double* moments;     // [nmodels][nT][DCEGM_SIM_NMOM]
    int32_t* ret_hist;   // [nmodels][nT+1]
    int32_t* sim_status; // [nmodels]
    double* pan_d;       // [nmodels][6][nsims*nT] or null
    int32_t* pan_i;      // [nmodels][2][nsims*nT]
    int32_t* pan_ret;    // [nmodels][nsims]
};

__device__ __forceinline__ size_t line_slot(const dcegm_layout& l, int m, int it, int iy, int id, int kind) {
    return ((((size_t)m * l.nT + it) * l.nstate + iy) * 2 + id) * 2 + kind;
}
__device__ __forceinline__ size_t state_slot(const dcegm_layout& l, int m, int it, int iy) {
    return ((size_t)m * l.nT + it) * l.nstate + iy;
}

// All kernel scratch lives in the dynamic shared memory and is addressed as integer offsets from
// this one symbol, so every access compiles to LDS/STS (pointer tables in registers or local
// memory made the compiler fall back to generic LD/ST).
extern __shared__ __align__(16) unsigned char dc_smem[];

// Where the scratch of a solver lives.  The persistent per-model kernels keep everything in shared
// memory (MemShared: offsets from dc_smem, every access an LDS/STS); the fine-grid solver
// (solver_f_big.cu: lines too long for one SM's shared memory) runs the same code on a per-model
// region of global memory (MemGlobal).  kTop = largest step of the fixed-trip knot search: lines hold
// fewer than 2*kTop points.
struct MemShared {
    static constexpr int kMaxWarps = 8;       // = kWarps: block-reduction slots in the envelope scratch
    static constexpr int kTop = 1024;
    static constexpr int kSegCapFixed = 192;  // max sections of one splitLine (more -> DCEGM_ST_SEG_OVERFLOW)
    static constexpr bool kLongLines = false;
    __device__ __forceinline__ unsigned char* mem() const { return dc_smem; }
};
struct MemGlobal {
    static constexpr int kMaxWarps = 32;
    static constexpr int kTop = 32768;
    static constexpr int kSegCapFixed = 0;    // room for one section per point: splitLine cannot overflow
    static constexpr bool kLongLines = true;  // tens of thousands of points per line: the envelope uses its
                                              // O(n log n) variants (merge sort, merged insertion, staged sections)
    unsigned char* g;
    __device__ __forceinline__ unsigned char* mem() const { return g; }
};

// Phase clocks (profiling builds only, -DDCEGM_PHASE_CLOCKS; see tools/phase_clocks.py): block-wide
// barrier, then thread 0 adds the cycles since the previous mark to slot i.  Compiles to nothing in
// the product build.
#ifdef DCEGM_PHASE_CLOCKS
constexpr int kPhaseSlots = 16;
__device__ __forceinline__ long long* dc_clk() {
    __shared__ long long s[kPhaseSlots + 1];
    return s;
}
#define DC_CLK(i)                                  \
    do {                                           \
        __syncthreads();                           \
        if (threadIdx.x == 0) {                    \
            long long c_ = clock64();              \
            dc_clk()[i] += c_ - dc_clk()[kPhaseSlots]; \
            dc_clk()[kPhaseSlots] = c_;            \
        }                                          \
    } while (0)
#else
#define DC_CLK(i) do { } while (0)
#endif

// ---------------------------------------------------------------------------------------------
// polyline primitives (line.jl:220-254).  Lines live in shared memory as SoA (x[], y[]).
// ---------------------------------------------------------------------------------------------
// number of knots <= q (searchsortedlast): fixed-trip, branch-free power-of-two descent.  kSearchTop
// is the largest step; lines hold fewer than 2*kSearchTop points (checked at upload).
constexpr int kSearchTop = MemShared::kTop;
template <int Top = kSearchTop>
__device__ __forceinline__ int count_le(const double* __restrict__ x, int n, double q) {
    int lo = 0;
#pragma unroll
    for (int step = Top; step >= 1; step >>= 1) {
        const int idx = lo + step - 1;
        const bool ok = idx < n && x[idx < n ? idx : 0] <= q;
        lo = ok ? lo + step : lo;
    }
    return lo;
}
template <int Top = kSearchTop>
__device__ __forceinline__ int left_knot(const double* __restrict__ x, int n, double q) {
    int i = count_le<Top>(x, n, q) - 1;
    i = i < 0 ? 0 : i;
    i = i > n - 2 ? n - 2 : i;
    return i;
}
// Gridded(Linear()) inside, extrapolate(.., Line()) outside; y only
__device__ __forceinline__ double interp_at(const double* __restrict__ x, const double* __restrict__ y, int n, int i,
                                            double q) {
    if (q < x[0]) {
        double d = x[1] - x[0];
        return y[0] + ((y[1] - y[0]) / d) * (q - x[0]);
    }
    if (q > x[n - 1]) {
        double d = x[n - 1] - x[n - 2];
        return y[n - 1] + ((y[n - 1] - y[n - 2]) / d) * (q - x[n - 1]);
    }
    double f = (q - x[i]) / (x[i + 1] - x[i]);
    return (1.0 - f) * y[i] + f * y[i + 1];
}
template <int Top = kSearchTop>
__device__ __forceinline__ double interp_y(const double* __restrict__ x, const double* __restrict__ y, int n, double q) {
    return interp_at(x, y, n, left_knot<Top>(x, n, q), q);
}
// ---------------------------------------------------------------------------------------------
// Bucket index over the knots of one line: T[b] = number of knots whose bucket is < b, for the
// monotone map bucket(q) = round(clamp((q - lo) * inv, 0, nb-1)).  Because the map is monotone in
// q (each step — subtraction, multiplication by a positive constant, clamp, rounding — is), all
// knots in lower buckets are <= q and all knots in higher buckets are > q, so count_le(q) is found
// by scanning only the knots of q's own bucket, starting at T[bucket(q)]: about two dependent
// shared-memory loads instead of the eleven of a binary search.  Exact for strictly increasing
// finite knots (lines that are not are flagged in the status word and searched by count_le).
// ---------------------------------------------------------------------------------------------
struct KnotIndex {
    int tab;        // byte offset in dc_smem of unsigned short T[nb + 1]
    double lo, inv;
    int topi;       // nb - 1
};
__device__ __forceinline__ int knot_bucket(const KnotIndex& K, double q) {
    // round to nearest by the magic-number addition, then clamp as an integer (two ALU min/max instead
    // of FP64 compare-select chains).  Monotone in q as long as |t| < 2^31 — knot_index_covers() checks
    // the queries of a period against that; NaN / Inf give bucket 0 (low word of the sum is 0).
    const double t = (q - K.lo) * K.inv;
    const int b = __double2loint(t + DC_MAGIC);
    return min(max(b, 0), K.topi);
}
// largest power of two nb such that nlines tables of nb+1 entries fit into `bytes`
__host__ __device__ inline int knot_index_buckets(int bytes, int nlines) {
#ifndef DC_KNOT_MAXNB
#define DC_KNOT_MAXNB 2048
#endif
    int nb = 16;
    while (nb < DC_KNOT_MAXNB && nlines * (2 * nb + 8) * 2 <= bytes) nb *= 2;  // stride (nb + 8) entries: 16-byte aligned
    return nb;
}
// lo / inv / top of the index over knots x[0..n) with nb buckets
__device__ __forceinline__ void knot_index_range(KnotIndex& K, const double* x, int n, int nb) {
    K.lo = x[0];
    K.inv = (double)nb / (x[n - 1] - x[0]);
    K.topi = nb - 1;
}
// true iff every query in [qmin, qmax] maps to a bucket number that fits the integer conversion of knot_bucket
__device__ __forceinline__ bool knot_index_covers(const KnotIndex& K, double qmin, double qmax) {
    const double t0 = (qmin - K.lo) * K.inv, t1 = (qmax - K.lo) * K.inv;
    return t0 > -1e9 && t1 < 1e9;  // false for NaN
}
// one knot's share of the table build (x strictly increasing, n >= 2): the buckets between the
// previous knot's and this knot's get this knot's index.  Block-level: every k in [0, n) once; the
// caller syncs afterwards.
__device__ __forceinline__ void knot_index_build_step_at(unsigned short* T, const KnotIndex& K, const double* __restrict__ x,
                                                          int n, int nb, int k) {
    const int bk = knot_bucket(K, x[k]);
    const int bp = k == 0 ? -1 : knot_bucket(K, x[k - 1]);
#pragma unroll 1
    for (int b = bp + 1; b <= bk; b++) T[b] = (unsigned short)k;
    if (k == n - 1) {
#pragma unroll 1
        for (int b = bk + 1; b <= nb; b++) T[b] = (unsigned short)n;
    }
}
__device__ __forceinline__ void knot_index_build_step(const KnotIndex& K, const double* __restrict__ x, int n, int nb, int k) {
    knot_index_build_step_at(reinterpret_cast<unsigned short*>(dc_smem + K.tab), K, x, n, nb, k);
}
// Block-level build of the tables of three lines at once, balanced whatever the spacing of the knots (the
// per-knot scheme above lets the thread that owns a knot behind a wide gap fill all the gap's buckets alone —
// 8 % of solve_f_kernel's time was spent waiting for it at the next barrier):
//   1. T[b] = n;  2. the first knot of every non-empty bucket writes its index;  3. suffix-minimum over the
// buckets (thread-local pass, warp shuffles, one exchange across warps) fills the empty ones.
// Gives T[b] = first knot whose bucket is >= b, the same table as knot_index_build_step.  All NT threads call it;
// the tables are valid after it returns.
template <int NT>
__device__ inline void knot_index_build3(unsigned short* T0, unsigned short* T1, unsigned short* T2, const KnotIndex& K0,
                                         const KnotIndex& K1, const KnotIndex& K2, const double* __restrict__ x0, int n0,
                                         const double* __restrict__ x1, int n1, const double* __restrict__ x2, int n2,
                                         int nb) {
    constexpr int NW = NT / 32;
    __shared__ int s_kmin[3][NW];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int b = tid; b <= nb; b += NT) {
        T0[b] = (unsigned short)n0;
        T1[b] = (unsigned short)n1;
        T2[b] = (unsigned short)n2;
    }
    __syncthreads();
    const int nmax = max(n0, max(n1, n2));
    for (int k = tid; k < nmax; k += NT) {
        if (k < n0) {
            const int bk = knot_bucket(K0, x0[k]);
            if (k == 0 || knot_bucket(K0, x0[k - 1]) != bk) T0[bk] = (unsigned short)k;
        }
        if (k < n1) {
            const int bk = knot_bucket(K1, x1[k]);
            if (k == 0 || knot_bucket(K1, x1[k - 1]) != bk) T1[bk] = (unsigned short)k;
        }
        if (k < n2) {
            const int bk = knot_bucket(K2, x2[k]);
            if (k == 0 || knot_bucket(K2, x2[k - 1]) != bk) T2[bk] = (unsigned short)k;
        }
    }
    __syncthreads();
    const int per = (nb + NT) / NT;  // entries 0..nb, `per` consecutive ones per thread
    const int b0 = tid * per;
    int mine[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
    for (int i = 0; i < per; i++) {
        const int b = b0 + i;
        if (b <= nb) {
            mine[0] = min(mine[0], (int)T0[b]);
            mine[1] = min(mine[1], (int)T1[b]);
            mine[2] = min(mine[2], (int)T2[b]);
        }
    }
    int carry[3];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        int v = mine[j];  // inclusive suffix minimum over the lanes of the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_down_sync(0xffffffffu, v, o);
            if (lane + o < 32) v = min(v, t);
        }
        if (lane == 0) s_kmin[j][wid] = v;
        const int nxt = __shfl_down_sync(0xffffffffu, v, 1);
        carry[j] = lane < 31 ? nxt : 0x7fffffff;  // minimum over the later lanes
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 3; j++)
        for (int w = wid + 1; w < NW; w++) carry[j] = min(carry[j], s_kmin[j][w]);
    for (int i = per - 1; i >= 0; i--) {
        const int b = b0 + i;
        if (b <= nb) {
            carry[0] = min(carry[0], (int)T0[b]);
            carry[1] = min(carry[1], (int)T1[b]);
            carry[2] = min(carry[2], (int)T2[b]);
            T0[b] = (unsigned short)carry[0];
            T1[b] = (unsigned short)carry[1];
            T2[b] = (unsigned short)carry[2];
        }
    }
    __syncthreads();
}

// number of knots <= q: all knots in lower buckets are <= q, so the scan starts at the first knot of
// q's own bucket.  (A variant with three independent probes instead of the dependent scan measured
// 4 % slower on the C3 batch: most buckets hold one knot.)
__device__ __forceinline__ int count_le_indexed_at(const unsigned short* __restrict__ T, const KnotIndex& K,
                                                    const double* __restrict__ x, int n, double q) {
    int i = T[knot_bucket(K, q)];
    while (i < n && x[i] <= q) i++;
    return i;
}
__device__ __forceinline__ int count_le_indexed(const KnotIndex& K, const double* __restrict__ x, int n, double q) {
    return count_le_indexed_at(reinterpret_cast<const unsigned short*>(dc_smem + K.tab), K, x, n, q);
}

// last statement of a per-model solver kernel: publish "this model's lines are in HBM" to the packing stream
__device__ __forceinline__ void signal_model_done(const DeviceBatch& B, int m) {
    if (!B.chunk_done) return;
    __syncthreads();  // every thread's stores of this model are issued
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(&B.chunk_done[B.model_chunk[m]], 1);
    }
}

// status bit of a knot pair x0, x1 (consecutive abscissae of a stored line) the previous period's interpolate()
// rejects: DCEGM_ST_DUP_KNOT for equal knots, DCEGM_ST_NONMONOTONE for a decreasing (or unordered, NaN) pair; in the
// first period — whose lines no step of the solve interpolates (dc_algo.jl:22-37 reads it+1 only) — the
// informational DCEGM_ST_NONMONO_FIRST for both.  0 when x1 > x0.
__device__ __forceinline__ int knot_pair_flag(double x0, double x1, bool first_period) {
    if (x1 > x0) return 0;
    if (first_period) return (int)DCEGM_ST_NONMONO_FIRST;
    return x1 == x0 ? (int)DCEGM_ST_DUP_KNOT : (int)DCEGM_ST_NONMONOTONE;
}
// status bits after which the knot vectors of a model are not strictly increasing (or not finite): the bucket index
// is not exact there and the EGM step takes its generic path (binary searches)
constexpr int kStatusNoIndex = (int)(DCEGM_ST_NONMONOTONE | DCEGM_ST_DUP_KNOT | DCEGM_ST_NONFINITE);

// whole-Point interpolation on a 2-point line (dc_algo.jl:446-478)
__device__ __forceinline__ double2 interp2pt(double2 a, double2 b, double q) {
    if (b.x < a.x) {  // sortx!(newleft)
        double2 t = a;
        a = b;
        b = t;
    }
    double2 r;
    double d = b.x - a.x;
    if (q < a.x) {
        r.x = a.x + ((b.x - a.x) / d) * (q - a.x);
        r.y = a.y + ((b.y - a.y) / d) * (q - a.x);
    } else if (q > b.x) {
        r.x = b.x + ((b.x - a.x) / d) * (q - b.x);
        r.y = b.y + ((b.y - a.y) / d) * (q - b.x);
    } else {
        // searchsortedlast clamps to the single interval
        double f = (q - a.x) / d;
        r.x = (1.0 - f) * a.x + f * b.x;
        r.y = (1.0 - f) * a.y + f * b.y;
    }
    return r;
}

// ---------------------------------------------------------------------------------------------
// scalar economics (utils.jl)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double u_work(double x, bool working, const DevModel& p) {  // utils.jl:40-46
    double dis = working ? p.alpha : 0.0;
    if (p.gamma_is_one) return dc_log(x) - dis;
    return p.oneover_oneminusgamma * (dc_pow(x, p.oneminusgamma) - 1.0) - dis;
}
__device__ __forceinline__ double u_plain(double x, const DevModel& p) {  // utils.jl:23-29
    if (p.gamma_is_one) return dc_log(x);
    return p.oneover_oneminusgamma * dc_pow(x, p.oneminusgamma);
}
__device__ __forceinline__ double up(double c, const DevModel& p) {  // utils.jl:72-78
    if (p.gamma_is_one) return 1.0 / c;
    return dc_pow(c, p.neg_gamma);
}
__device__ __forceinline__ double iup(double u, const DevModel& p) {  // utils.jl:102-108
    if (p.gamma_is_one) return 1.0 / u;
    return dc_pow(u, p.neg_oneover_gamma);
}

// ---------------------------------------------------------------------------------------------
// block helpers (kThreads threads)
// ---------------------------------------------------------------------------------------------
// The cooperative routines below and worker_envelope (envelope.cuh) run on a THREAD GROUP of NT threads (their template
// parameter): the whole CTA by default.  A .cu file that defines DC_GROUPED 1 before including this header runs them on
// groups of NT consecutive threads of a wider CTA (NT a multiple of 32, at most 15 groups), every group on its own named
// barrier, so that several groups work through independent envelopes at their own pace (solver_g.cu: one income state
// per group).
#ifndef DC_GROUPED
#define DC_GROUPED 0
#endif
template <int NT>
__device__ __forceinline__ int dc_gtid() {
    return DC_GROUPED ? (int)(threadIdx.x % NT) : (int)threadIdx.x;
}
template <int NT>
__device__ __forceinline__ void dc_gsync() {
#if DC_GROUPED
    asm volatile("bar.sync %0, %1;" ::"r"(1 + (int)(threadIdx.x / NT)), "n"(NT) : "memory");
#else
    __syncthreads();
#endif
}
// exclusive prefix sum of data[0..n) in place; returns the total.  red: (NT / 32)+1 ints of scratch.
template <int NT = kThreads>
__device__ inline int block_scan_excl(int* data, int n, int* red) {
    const int tid = dc_gtid<NT>(), lane = tid & 31, wid = tid >> 5;
    const int per = (n + NT - 1) / NT;
    const int b = tid * per;
    int local = 0;
#pragma unroll (NT > 256 ? 4 : 1)
    for (int i = 0; i < per; i++) {
        int k = b + i;
        if (k < n) local += data[k];
    }
    int inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) red[wid] = inc;
    dc_gsync<NT>();
    if (wid == 0) {
        int v = lane < (NT / 32) ? red[lane] : 0;
        int vi = v;
#pragma unroll
        for (int o = 1; o < (NT / 32); o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, vi, o);
            if (lane >= o) vi += t;
        }
        if (lane < (NT / 32)) red[lane] = vi - v;
        if (lane == (NT / 32) - 1) red[(NT / 32)] = vi;
    }
    dc_gsync<NT>();
    int run = red[wid] + inc - local;
    const int total = red[(NT / 32)];
#pragma unroll (NT > 256 ? 4 : 1)
    for (int i = 0; i < per; i++) {
        int k = b + i;
        if (k < n) {
            int v = data[k];
            data[k] = run;
            run += v;
        }
    }
    dc_gsync<NT>();
    return total;
}

template <int NT = kThreads>
__device__ inline double block_min(double v, double* red) {
    const int lane = dc_gtid<NT>() & 31, wid = dc_gtid<NT>() >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane == 0) red[wid] = v;
    dc_gsync<NT>();
    double r = red[0];
#pragma unroll
    for (int w = 1; w < (NT / 32); w++) r = fmin(r, red[w]);
    dc_gsync<NT>();
    return r;
}
template <int NT = kThreads>
__device__ inline int block_sum(int v, int* red) {
    const int lane = dc_gtid<NT>() & 31, wid = dc_gtid<NT>() >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[wid] = v;
    dc_gsync<NT>();
    int r = 0;
#pragma unroll
    for (int w = 0; w < (NT / 32); w++) r += red[w];
    dc_gsync<NT>();
    return r;
}
template <int NT = kThreads>
__device__ inline int block_min_int(int v, int* red) {
    const int lane = dc_gtid<NT>() & 31, wid = dc_gtid<NT>() >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane == 0) red[wid] = v;
    dc_gsync<NT>();
    int r = red[0];
#pragma unroll
    for (int w = 1; w < (NT / 32); w++) r = min(r, red[w]);
    dc_gsync<NT>();
    return r;
}

}  // namespace dcegm
