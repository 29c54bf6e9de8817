// solver_f_big.cu — fine-grid solver of the retirement model: grids whose polylines do not fit into one
// SM's shared memory (BASELINE config "fine-grid retirement model": na = 20 000, nT = 50, ny = 32).
//
// Same algorithm, same arithmetic and the same output buffers as solve_f_kernel (solver_f.cu) —
// dc_EGM!(m::FModel, p), src/dc_algo.jl:8-105 — but the work of one period is spread over the whole GPU:
//   big_terminal_kernel  one CTA per model      terminal lines (dc_algo.jl:14-20) + knot index
//   big_egm_kernel       CTA per (model, 256 savings grid points): the EGM step (egm_point.inc: the very
//                        code solve_f_kernel runs) reading the next-period lines of the model from its
//                        region of global memory (L2-resident: 4 lines x 16 B x na per model)
//   big_env_kernel       one CTA per model: secondary upper envelope + policy repair (envelope.cuh,
//                        instantiated on global-memory scratch), credit-constraint point
//   big_post_kernel      CTA per (model, 256-point / 256-bucket tile): everything of the period's tail that is per-point
//                        work — the one store of the finished lines, the retiree line's checks, and the knot index of
//                        the new lines (one binary search per bucket) — which took a third of big_env_kernel's time
//                        when its single CTA per model did it (profiles/r1g_big_env_by_source_line.txt)
// Three launches per period; every period's lines are written to the solution buffers once, the working
// copies (SoA, two period sets) live in a per-model scratch region.
#include <math_constants.h>

#include <cstdlib>
#include "device_common.cuh"
#include "egm_node.cuh"
#include "envelope.cuh"

namespace dcegm {

// threads of big_env_kernel: the envelope of one model is one CTA's work, its passes over the ~22 000 points
// of a C4 line are bound by memory latency, so the CTA is as wide as the register file allows
#ifndef DCEGM_BIG_ENV_THREADS
#define DCEGM_BIG_ENV_THREADS 1024
#endif
constexpr int kEnvThreads = DCEGM_BIG_ENV_THREADS;
#ifndef DCEGM_BIG_EGM_MINBLOCKS
#define DCEGM_BIG_EGM_MINBLOCKS 3
#endif

// per-model state that survives between the launches of one solve
struct BigState {
    int n_wc[2], n_wv[2], n_r[2];  // by line set
    double vb_w[2], vb_r[2];
    KnotIndex Kr, Kc, Kv;           // index over the lines of the set built last
    int idx_ok;                     // the three tables were built (lines of at least two points)
    int status;
    int now;                        // status bits raised by the lines of the period built last
    double ev0[2];
    int sum_envp, sum_seg, sum_is, sum_rm, sum_maxlen;
    long long sum_pts;
};

// byte map of one model's scratch region
struct BigPlan {
    int cap8;          // bytes of one double[cap] array
    EnvScratchG env;   // envelope scratch behind the two line sets
    int nbuck, tab_entries;  // buckets of one knot-index table, entries of one table incl. padding
    int o_tab, o_state;
    size_t bytes;
    __host__ __device__ BigPlan(int cap, int isec_cap) {
        cap8 = EnvScratchG::align16(cap * 8);
        env.g = nullptr;
        env.off = 14 * cap8;
        env.cap = cap;
        env.isec_cap = isec_cap;
        nbuck = 16;
        while (nbuck < cap && nbuck < 32768) nbuck *= 2;
        tab_entries = nbuck + 8;
        o_tab = env.off + env.bytes();
        o_state = o_tab + EnvScratchG::align16(3 * tab_entries * 2);
        bytes = (size_t)o_state + EnvScratchG::align16((int)sizeof(BigState));
        bytes = (bytes + 255) / 256 * 256;
    }
    __device__ __forceinline__ unsigned short* tab(unsigned char* g, int j) const {
        return reinterpret_cast<unsigned short*>(g + o_tab) + j * tab_entries;
    }
    __device__ __forceinline__ BigState* state(unsigned char* g) const { return reinterpret_cast<BigState*>(g + o_state); }
};

size_t solver_f_big_scratch_bytes(int cap, int isec_cap) { return BigPlan(cap, isec_cap).bytes; }

__device__ __forceinline__ LineSetG big_set(const BigPlan& P, unsigned char* g, const BigState* st, int which) {
    LineSetG L;
    L.g = g;
    L.cap8 = P.cap8;
    L.base = which * 7 * P.cap8;
    L.n_wc = st->n_wc[which];
    L.n_wv = st->n_wv[which];
    L.n_r = st->n_r[which];
    L.vb_w = st->vb_w[which];
    L.vb_r = st->vb_r[which];
    return L;
}

// block-level: knot index over the three knot vectors of L (the lines the next EGM launch reads)
template <int NT>
__device__ inline void big_build_index(const BigPlan& P, unsigned char* g, BigState* st, const LineSetG& L, int status) {
    const int tid = threadIdx.x;
    const bool ok = !(status & kStatusNoIndex) && L.n_wc >= 2 && L.n_wv >= 2 && L.n_r >= 2;
    KnotIndex Kr, Kc, Kv;
    Kr.tab = Kc.tab = Kv.tab = 0;
    Kr.lo = Kc.lo = Kv.lo = Kr.inv = Kc.inv = Kv.inv = 0.0;
    Kr.topi = Kc.topi = Kv.topi = 0;
    if (ok) {
        const int nb = P.nbuck;
        knot_index_range(Kr, L.r_x(), L.n_r, nb);
        knot_index_range(Kc, L.wc_x(), L.n_wc, nb);
        knot_index_range(Kv, L.wv_x(), L.n_wv, nb);
        knot_index_build3<NT>(P.tab(g, 0), P.tab(g, 1), P.tab(g, 2), Kr, Kc, Kv, L.r_x(), L.n_r, L.wc_x(), L.n_wc, L.wv_x(),
                              L.n_wv, nb);
    }
    if (tid == 0) {
        st->Kr = Kr;
        st->Kc = Kc;
        st->Kv = Kv;
        st->idx_ok = ok ? 1 : 0;
    }
}

// block-level: per-model summary record (payload of the multi-GPU allgather), same arithmetic and summation
// order as the tail of solve_f_kernel
__device__ inline void big_summary(const DeviceBatch& B, int m, const DevModel& p, const LineSetG& L0, const BigState* st,
                                   int status) {
    const int tid = threadIdx.x;
    const int nT = p.nT;
    constexpr int Top = MemGlobal::kTop;
    // the first kThreads threads sum in the order of solve_f_kernel's 256-thread CTA, whatever this CTA's width
    double cs_c = 0.0, cs_v = 0.0;
    __shared__ double s_cs[2][kWarps];
    if (tid < kThreads) {
        for (int k = tid; k < L0.n_wc; k += kThreads) cs_c += L0.wc_y()[k];
        for (int k = tid; k < L0.n_r; k += kThreads) cs_c += L0.rc_y()[k];
        for (int k = tid; k < L0.n_wv; k += kThreads) cs_v += isfinite(L0.wv_y()[k]) ? L0.wv_y()[k] : 0.0;
        for (int k = tid; k < L0.n_r; k += kThreads) cs_v += isfinite(L0.rv_y()[k]) ? L0.rv_y()[k] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cs_c += __shfl_xor_sync(0xffffffffu, cs_c, o);
            cs_v += __shfl_xor_sync(0xffffffffu, cs_v, o);
        }
        if ((tid & 31) == 0) {
            s_cs[0][tid >> 5] = cs_c;
            s_cs[1][tid >> 5] = cs_v;
        }
    }
    __syncthreads();
    if (tid == 0) {
        dcegm_summary sm;
        sm.status = status;
        sm.n_env_periods = st->sum_envp;
        sm.tot_segments = st->sum_seg;
        sm.tot_isects = st->sum_is;
        sm.tot_removed = st->sum_rm;
        sm.max_len = st->sum_maxlen;
        sm.tot_points = st->sum_pts;
        const double fr[DCEGM_SUMMARY_NPROBE] = {0.02, 0.05, 0.1, 0.2, 0.3, 0.5, 0.7, 0.9};
        for (int q = 0; q < DCEGM_SUMMARY_NPROBE; q++) {
            const double mq = p.a_low + fr[q] * (p.a_high - p.a_low);
            double c = interp_y<Top>(L0.wc_x(), L0.wc_y(), L0.n_wc, mq);
            c = c < p.cfloor ? p.cfloor : c;
            sm.c1_probe[q] = c;
            sm.v1_probe[q] = nT >= 2 ? vfun_dev<Top>(true, false, c, mq, L0.wv_x(), L0.wv_y(), L0.n_wv, L0.vb_w, p) : 0.0;
        }
        double a = 0.0, b = 0.0;
        for (int wi = 0; wi < kWarps; wi++) {
            a += s_cs[0][wi];
            b += s_cs[1][wi];
        }
        sm.checksum_c = a;
        sm.checksum_v = b;
        B.summaries[m] = sm;
        B.status[m] = status;
    }
}

// ---- terminal period (dc_algo.jl:14-20) -------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) big_terminal_kernel(DeviceBatch B, unsigned char* scratch, size_t stride) {
    const int m = B.first + blockIdx.x;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    if (tid == 0) s_p = B.models[m];
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, nT = p.nT;
    const BigPlan P(B.lay.cap, B.lay.isec_cap);
    unsigned char* g = scratch + (size_t)blockIdx.x * stride;
    BigState* st = P.state(g);
    EnvScratchG S = P.env;
    S.g = g;
    const int t = nT - 1, w = t & 1;
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
    LineSetG C;
    C.g = g;
    C.cap8 = P.cap8;
    C.base = w * 7 * P.cap8;
    if (tid == 0) {
        C.wc_x()[0] = p.a_lowT; C.wc_y()[0] = 0.0; C.wc_x()[1] = p.a_high; C.wc_y()[1] = p.a_high;
        C.r_x()[0] = p.a_lowT;  C.rc_y()[0] = 0.0; C.r_x()[1] = p.a_high;  C.rc_y()[1] = p.a_high;
    }
    int nv = 2;
    if (p.nu > 0) {
        nv = na;
        for (int k = tid; k < na; k += kThreads) {
            double b = p.nu * u_plain(avec[k] + p.bbar, p);
            C.wv_x()[k] = avec[k];
            C.wv_y()[k] = b - p.alpha;  // bequest(a, true, p)
        }
    } else if (tid == 0) {
        C.wv_x()[0] = p.a_lowT; C.wv_y()[0] = 0.0; C.wv_x()[1] = p.a_high; C.wv_y()[1] = kNaN;
        C.rv_y()[0] = 0.0; C.rv_y()[1] = kNaN;
    }
    C.n_wc = 2; C.n_wv = nv; C.n_r = 2;
    C.vb_w = 0.0; C.vb_r = 0.0;
    __syncthreads();
    store_line(B, m, t, 0, 0, DCEGM_KIND_C, C.wc_x(), C.wc_y(), 2);
    store_line(B, m, t, 0, 1, DCEGM_KIND_C, C.r_x(), C.rc_y(), 2);
    if (p.nu > 0) {
        for (int k = tid; k < na; k += kThreads) {
            S.Dx()[k] = avec[k];
            S.Dv()[k] = p.nu * u_plain(avec[k] + p.bbar, p) - 0.0;
        }
        __syncthreads();
        store_line(B, m, t, 0, 0, DCEGM_KIND_V, C.wv_x(), C.wv_y(), na);
        store_line(B, m, t, 0, 1, DCEGM_KIND_V, S.Dx(), S.Dv(), na);
    } else {
        store_line(B, m, t, 0, 0, DCEGM_KIND_V, C.wv_x(), C.wv_y(), 2);
        store_line(B, m, t, 0, 1, DCEGM_KIND_V, C.r_x(), C.rv_y(), 2);
    }
    if (tid == 0) {
        B.vbound[state_slot(B.lay, m, t, 0) * 2 + 0] = 0.0;
        B.vbound[state_slot(B.lay, m, t, 0) * 2 + 1] = 0.0;
        dcegm_env_meta z = {0, 0, 0, 0};
        B.meta[state_slot(B.lay, m, t, 0)] = z;
        st->n_wc[w] = 2; st->n_wv[w] = nv; st->n_r[w] = 2;
        st->vb_w[w] = 0.0; st->vb_r[w] = 0.0;
        st->status = 0;
        st->now = 0;
        st->ev0[0] = st->ev0[1] = 0.0;
        st->sum_envp = st->sum_seg = st->sum_is = st->sum_rm = st->sum_maxlen = 0;
        st->sum_pts = 4 + 2 * nv;
    }
    big_build_index<kThreads>(P, g, st, C, 0);
    if (nT == 1) {
        __syncthreads();
        big_summary(B, m, p, C, st, 0);
    }
}

// ---- EGM step of period t: thread per savings grid point ------------------------------------------------
__global__ void __launch_bounds__(kThreads, DCEGM_BIG_EGM_MINBLOCKS) big_egm_kernel(DeviceBatch B, unsigned char* scratch, size_t stride, int t) {
    const int lm = blockIdx.y;  // model within this launch
    const int m = B.first + lm;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    __shared__ double s_off[2 * kMaxNy], s_wgt[kMaxNy];
    if (tid == 0) s_p = B.models[m];
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const BigPlan P(B.lay.cap, B.lay.isec_cap);
    unsigned char* g = scratch + (size_t)lm * stride;
    BigState* st = P.state(g);
    EnvScratchG S = P.env;
    S.g = g;
    const LineSetG N = big_set(P, g, st, (t + 1) & 1);
    LineSetG C;
    C.g = g;
    C.cap8 = P.cap8;
    C.base = (t & 1) * 7 * P.cap8;

    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny;
    const double* __restrict__ offm = B.off + (size_t)m * (nT - 1) * 2 * ny;
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const bool can_skip = p.cfloor > 0.0;
    const bool model_fast = isfinite(p.neg_gamma) && isfinite(p.oneminusgamma) && isfinite(p.neg_oneover_gamma) &&
                            isfinite(p.oneover_oneminusgamma) && dc_is_norm_pos(p.cfloor) && p.lambda > 1e-100 && p.lambda < 1e100;
    if (tid < ny) s_wgt[tid] = wgt[tid];
    if (tid < 2 * ny) s_off[tid] = offm[(size_t)t * 2 * ny + tid];
    __syncthreads();

    const bool next_terminal = (t + 1 == nT - 1);
    const bool beq = (t == nT - 2) && p.nu > 0;
    const double scale = beq ? p.betaRnu : p.betaR;
    // the bucket index (and with it the straight-line node) needs the next-period lines strictly increasing and finite
    bool fast = model_fast && !beq && st->idx_ok && !(st->now & kStatusNoIndex);
    const KnotIndex Kr = st->Kr, Kc = st->Kc, Kv = st->Kv;
    const unsigned short* __restrict__ Tr = P.tab(g, 0);
    const unsigned short* __restrict__ Tc = P.tab(g, 1);
    const unsigned short* __restrict__ Tv = P.tab(g, 2);
    bool same_nodes = true;
    for (int iy = 1; iy < ny; iy++) same_nodes &= s_off[ny + iy] == s_off[ny];
    if (fast) {
        double omin = s_off[0], omax = s_off[0];
        for (int k = 1; k < 2 * ny; k++) {
            omin = s_off[k] < omin ? s_off[k] : omin;
            omax = s_off[k] > omax ? s_off[k] : omax;
        }
        const double ra0 = avec[0] * p.R, ra1 = avec[na - 1] * p.R;
        const double qmin = (ra0 < ra1 ? ra0 : ra1) + omin, qmax = (ra0 < ra1 ? ra1 : ra0) + omax;
        fast = knot_index_covers(Kr, qmin, qmax) && knot_index_covers(Kc, qmin, qmax) && knot_index_covers(Kv, qmin, qmax);
    }
    const double rx1 = N.r_x()[1], wvx1 = N.wv_x()[1];
    const double* __restrict__ rX = N.r_x();
    const double* __restrict__ rC = N.rc_y();
    const double* __restrict__ rV = N.rv_y();
    const double* __restrict__ cX = N.wc_x();
    const double* __restrict__ cY = N.wc_y();
    const double* __restrict__ vX = N.wv_x();
    const double* __restrict__ vY = N.wv_y();
    const int n_r = N.n_r, n_wc = N.n_wc, n_wv = N.n_wv;
    const double bvb_r = p.beta * N.vb_r, bvb_w = p.beta * N.vb_w;
    const bool fast_r = fast && p.delta == 0.0 && can_skip;
    const bool same = same_nodes && fast_r;
    double* s_ev0 = st->ev0;
    const int ia = blockIdx.x * kThreads + tid;
    if (ia < na) {
        const double a = avec[ia];
#define EGM_TAB(K) ((&K == &Kr) ? Tr : (&K == &Kc) ? Tc : Tv)
#define EGM_COUNT_LE(K, x, n, q) count_le_indexed_at(EGM_TAB(K), K, x, n, q)
#define EGM_LINESET LineSetG
#include "egm_point.inc"
#undef EGM_COUNT_LE
#undef EGM_LINESET
#undef EGM_TAB
    }
}

#ifdef DCEGM_PHASE_CLOCKS
__device__ unsigned long long dc_phase_total_big[kPhaseSlots];
extern "C" int dcegm_debug_phase_clocks_big(unsigned long long* out, int reset) {
    if (out && cudaMemcpyFromSymbol(out, dc_phase_total_big, sizeof(unsigned long long) * kPhaseSlots) != cudaSuccess) return -2;
    if (reset) {
        unsigned long long z[kPhaseSlots] = {0};
        if (cudaMemcpyToSymbol(dc_phase_total_big, z, sizeof z) != cudaSuccess) return -2;
    }
    return 0;
}
#define BIG_CLK_BEGIN()                                           \
    if (threadIdx.x == 0) {                                       \
        for (int i = 0; i < kPhaseSlots; i++) dc_clk()[i] = 0;    \
        dc_clk()[kPhaseSlots] = clock64();                        \
    }
#define BIG_CLK_END()                                             \
    do {                                                          \
        DC_CLK(14);                                               \
        if (threadIdx.x == 0)                                     \
            for (int i = 0; i < kPhaseSlots; i++) atomicAdd(&dc_phase_total_big[i], (unsigned long long)dc_clk()[i]); \
    } while (0)
#else
#define BIG_CLK_BEGIN()
#define BIG_CLK_END() do { } while (0)
#endif

// ---- envelope, tail and store of period t: one CTA per model ------------------------------------------------
__global__ void __launch_bounds__(kEnvThreads) big_env_kernel(DeviceBatch B, unsigned char* scratch, size_t stride, int t) {
    const int m = B.first + blockIdx.x;
    const int tid = threadIdx.x;
    BIG_CLK_BEGIN();
    __shared__ DevModel s_p;
    __shared__ int s_n[2];
    __shared__ int s_status;  // status word of the model so far
    if (tid == 0) s_p = B.models[m];
    const BigPlan P(B.lay.cap, B.lay.isec_cap);
    unsigned char* g = scratch + (size_t)blockIdx.x * stride;
    BigState* st = P.state(g);
    if (tid == 0) s_status = st->status;
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na;
    EnvScratchG S = P.env;
    S.g = g;
    const int w = t & 1;
    LineSetG C;
    C.g = g;
    C.cap8 = P.cap8;
    C.base = w * 7 * P.cap8;
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const double a0 = avec[0];
    const double ev0w = st->ev0[0], ev0r = st->ev0[1];

    if (B.dirty) {  // runfdemo: vdirty / cdirty (dc_algo.jl:79-82)
        double2* d = B.dirty + line_slot(B.lay, m, t, 0, 0, 0) * (size_t)na;
        for (int k = tid; k < na; k += kEnvThreads) {
            d[k] = make_double2(S.Dx()[koff + k], S.Dc()[koff + k]);
            d[(size_t)na + k] = make_double2(S.Dx()[koff + k], S.Dv()[koff + k]);
            d[(size_t)2 * na + k] = make_double2(C.r_x()[1 + k], C.rc_y()[1 + k]);
            d[(size_t)3 * na + k] = make_double2(C.r_x()[1 + k], C.rv_y()[1 + k]);
        }
    }

    // worker: secondary envelope + policy repair
    const size_t ss = state_slot(B.lay, m, t, 0);
    const bool first_period = t == 0;  // see solve_f_kernel
    worker_envelope<EnvScratchG, kEnvThreads>(S, koff, na, a0, ev0w, p, C.wv_x(), C.wv_y(), &s_n[0], C.wc_x(), C.wc_y(), &s_n[1], B.meta + ss,
                    B.isects + ss * (size_t)B.lay.isec_cap, B.removed + ss * (size_t)B.lay.rem_words, B.lay.rem_words, first_period);
    C.n_wv = s_n[0];
    C.n_wc = s_n[1];
    C.vb_w = ev0w;
    DC_CLK(12);

    // retiree: credit-constraint point (dc_algo.jl:90-101); its checks, all stores and the knot index are
    // big_post_kernel's (per-point work, spread over the GPU)
    __syncthreads();
    if (tid == 0) {
        C.r_x()[0] = a0;
        C.rc_y()[0] = 0.0;
        C.rv_y()[0] = ev0r;
        const int wf = S.cnt()[2];  // bits raised by the worker's envelope
        B.vbound[ss * 2 + 0] = ev0w;
        B.vbound[ss * 2 + 1] = ev0r;
        const dcegm_env_meta mt = B.meta[ss];
        const int n_r = na + 1;
        st->sum_envp += mt.n_removed > 0;
        st->sum_seg += mt.n_segments;
        st->sum_is += mt.n_isects;
        st->sum_rm += mt.n_removed;
        int ml = C.n_wc > C.n_wv ? C.n_wc : C.n_wv;
        ml = ml > n_r ? ml : n_r;
        st->sum_maxlen = st->sum_maxlen > ml ? st->sum_maxlen : ml;
        st->sum_pts += C.n_wc + C.n_wv + 2 * n_r;
        st->n_wc[w] = C.n_wc;
        st->n_wv[w] = C.n_wv;
        st->n_r[w] = n_r;
        st->vb_w[w] = C.vb_w;
        st->vb_r[w] = ev0r;
        st->now = wf;
        st->status = s_status | wf;
    }
    BIG_CLK_END();
}

// ---- the per-point tail of period t: stores, retiree checks, knot index -------------------------------------------
__global__ void __launch_bounds__(kThreads) big_post_kernel(DeviceBatch B, unsigned char* scratch, size_t stride, int t) {
    const int lm = blockIdx.y;
    const int m = B.first + lm;
    const int tid = threadIdx.x;
    const BigPlan P(B.lay.cap, B.lay.isec_cap);
    unsigned char* g = scratch + (size_t)lm * stride;
    BigState* st = P.state(g);
    const int w = t & 1;
    const LineSetG C = big_set(P, g, st, w);
    const int n_wc = C.n_wc, n_wv = C.n_wv, n_r = C.n_r;
    const bool first_period = t == 0;
    const int k = blockIdx.x * kThreads + tid;
    // ---- the one store of the finished lines (interleaved x, y) ----
    {
        const size_t s0 = line_slot(B.lay, m, t, 0, 0, 0);  // slots: worker c, worker v, retiree c, retiree v
        if (k == 0) {
            B.lens[s0 + 0] = n_wc;
            B.lens[s0 + 1] = n_wv;
            B.lens[s0 + 2] = n_r;
            B.lens[s0 + 3] = n_r;
        }
        if (B.lines) {
            const int cap = B.lay.cap;
            double2* d = B.lines + s0 * (size_t)cap;
            if (k < n_wc && k < cap) d[k] = make_double2(C.wc_x()[k], C.wc_y()[k]);
            if (k < n_wv && k < cap) d[(size_t)cap + k] = make_double2(C.wv_x()[k], C.wv_y()[k]);
            if (k < n_r && k < cap) {
                const double x = C.r_x()[k];
                d[(size_t)2 * cap + k] = make_double2(x, C.rc_y()[k]);
                d[(size_t)3 * cap + k] = make_double2(x, C.rv_y()[k]);
            }
        }
    }
    // ---- retiree line: strictly increasing, finite ----
    {
        int flag = 0;
        if (k < n_r) {
            if (k + 1 < n_r) flag |= knot_pair_flag(C.r_x()[k], C.r_x()[k + 1], first_period);
            if (!isfinite(C.r_x()[k]) || !isfinite(C.rc_y()[k]) || !isfinite(C.rv_y()[k])) flag |= DCEGM_ST_NONFINITE;
        }
        flag = __reduce_or_sync(0xffffffffu, flag);
        if (flag && (tid & 31) == 0) {
            atomicOr(&st->status, flag);
            atomicOr(&st->now, flag);
        }
    }
    if (first_period) return;  // period 1 is read by nobody: no index
    // ---- knot index of the three knot vectors: entry b = first knot whose bucket is >= b, one search per entry ----
    const bool ok = n_wc >= 2 && n_wv >= 2 && n_r >= 2;
    const int nb = P.nbuck;
    KnotIndex K[3];
    const double* X[3] = {C.r_x(), C.wc_x(), C.wv_x()};
    const int N[3] = {n_r, n_wc, n_wv};
#pragma unroll
    for (int j = 0; j < 3; j++) {
        K[j].tab = 0;
        K[j].lo = K[j].inv = 0.0;
        K[j].topi = 0;
        if (ok) knot_index_range(K[j], X[j], N[j], nb);
    }
    if (k == 0) {
        st->Kr = K[0];
        st->Kc = K[1];
        st->Kv = K[2];
        st->idx_ok = ok ? 1 : 0;
    }
    if (ok && k <= nb) {
#pragma unroll
        for (int j = 0; j < 3; j++) {
            int lo = 0, hi = N[j];  // knot_bucket is non-decreasing along an increasing knot vector
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (knot_bucket(K[j], X[j][mid]) >= k)
                    hi = mid;
                else
                    lo = mid + 1;
            }
            P.tab(g, j)[k] = (unsigned short)lo;
        }
    }
}

// ---- per-model summary after the first period --------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) big_summary_kernel(DeviceBatch B, unsigned char* scratch, size_t stride) {
    const int m = B.first + blockIdx.x;
    __shared__ DevModel s_p;
    if (threadIdx.x == 0) s_p = B.models[m];
    __syncthreads();
    const BigPlan P(B.lay.cap, B.lay.isec_cap);
    unsigned char* g = scratch + (size_t)blockIdx.x * stride;
    BigState* st = P.state(g);
    const LineSetG C = big_set(P, g, st, 0);  // period 1 (t = 0) lives in set 0
    big_summary(B, m, s_p, C, st, st->status);
}

// Launches the whole backward recursion for models [first, first+count).  `scratch` holds `count` regions of
// solver_f_big_scratch_bytes().  All models of a batch share nT.  Returns the number of launches in *nlaunch.
// The models of a launch are split into parts that run on streams of their own: a model's envelope CTA fills an SM
// (1024 threads x 64 registers), the envelope is latency-bound, and a batch that is not a multiple of the SM count
// leaves SMs idle in its last wave — the other parts' EGM steps use that time (models are independent; results do not
// depend on the split).  Measured on C4 x256: 407 ms on one stream, 369 / 360 / 341 / 344 ms in 2 / 3 / 4 / 6 parts.
cudaError_t launch_solve_f_big(const DeviceBatch& B0, int first, int count, unsigned char* scratch, cudaStream_t stream,
                               int* nlaunch) {
    constexpr int kMaxParts = 8;
    const size_t stride = BigPlan(B0.lay.cap, B0.lay.isec_cap).bytes;
    const int nT = B0.lay.nT, na = B0.lay.na;
    const int nbuck_plus1 = BigPlan(B0.lay.cap, B0.lay.isec_cap).nbuck + 1;
    static const int parts_wanted = [] {
        const char* e = getenv("DCEGM_BIG_PARTS");  // 1 = one stream
        const int v = e ? atoi(e) : 4;
        return v < 1 ? 1 : (v > kMaxParts ? kMaxParts : v);
    }();
    const int np = count / 16 < 1 ? 1 : (count / 16 < parts_wanted ? count / 16 : parts_wanted);  // parts of >= 16 models
    cudaStream_t st[kMaxParts] = {stream};
    cudaEvent_t fork = nullptr, join[kMaxParts] = {};
    if (np > 1) {
        cudaError_t e = cudaEventCreateWithFlags(&fork, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
        cudaEventRecord(fork, stream);
        for (int h = 1; h < np; h++) {
            if ((e = cudaStreamCreateWithFlags(&st[h], cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&join[h], cudaEventDisableTiming)) != cudaSuccess) return e;
            cudaStreamWaitEvent(st[h], fork, 0);
        }
    }
    DeviceBatch Bh[kMaxParts];
    int cnt[kMaxParts], beg[kMaxParts];
    for (int h = 0; h < np; h++) {
        beg[h] = (int)((long long)count * h / np);
        cnt[h] = (int)((long long)count * (h + 1) / np) - beg[h];
        Bh[h] = B0;
        Bh[h].first = first + beg[h];
    }
    int nl = 0;
    for (int h = 0; h < np; h++) {
        big_terminal_kernel<<<cnt[h], kThreads, 0, st[h]>>>(Bh[h], scratch + (size_t)beg[h] * stride, stride);
        nl++;
    }
    for (int t = nT - 2; t >= 0; t--) {
        for (int h = 0; h < np; h++) {
            const dim3 ggrid((na + kThreads - 1) / kThreads, cnt[h]);
            unsigned char* scr = scratch + (size_t)beg[h] * stride;
            big_egm_kernel<<<ggrid, kThreads, 0, st[h]>>>(Bh[h], scr, stride, t);
            big_env_kernel<<<cnt[h], kEnvThreads, 0, st[h]>>>(Bh[h], scr, stride, t);
            const int ptiles = (B0.lay.cap > nbuck_plus1 ? B0.lay.cap : nbuck_plus1);
            big_post_kernel<<<dim3((ptiles + kThreads - 1) / kThreads, cnt[h]), kThreads, 0, st[h]>>>(Bh[h], scr, stride, t);
            nl += 3;
        }
    }
    for (int h = 0; h < np; h++) {
        big_summary_kernel<<<cnt[h], kThreads, 0, st[h]>>>(Bh[h], scratch + (size_t)beg[h] * stride, stride);
        nl++;
    }
    if (np > 1) {
        for (int h = 1; h < np; h++) {
            cudaEventRecord(join[h], st[h]);
            cudaStreamWaitEvent(stream, join[h], 0);
            cudaEventDestroy(join[h]);
            cudaStreamDestroy(st[h]);  // released when its work is done
        }
        cudaEventDestroy(fork);
    }
    if (nlaunch) *nlaunch = nl;
    return cudaGetLastError();
}

}  // namespace dcegm
