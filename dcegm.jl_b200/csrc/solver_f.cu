// solver_f.cu — persistent per-model kernel for the retirement model with iid income.
//
// Replaces dc_EGM!(m::FModel, p) (src/dc_algo.jl:8-105): one CTA owns one model
// parameterisation and walks the periods backwards; the four next-period polylines
// (policy/value x worker/retiree) stay in shared memory, every finished line is written to
// HBM exactly once.  Per period:
//   1. EGM step (dc_algo.jl:25-78): thread per (current choice, savings grid point); loop over
//      the income nodes and the two future choices: interpolate next-period consumption and
//      value at next-period cash-on-hand (interp, vfun), taste-shock choice probability and
//      log-sum (ccp, logsum), quadrature sums in ascending node order, Euler inversion (iup).
//   2. worker: secondary upper envelope + policy repair (envelope.cuh).
//   3. credit-constraint point, checks, store.
#include <math_constants.h>

#define DC_OUTLINE_MATH  // dc_math.h: out-of-line log/exp cores (instruction-cache footprint of the hot loop)

#include "device_common.cuh"
#include "egm_node.cuh"
#include "envelope.cuh"

namespace dcegm {

// shared-memory map of solve_f_kernel: 2 sets x 7 double[cap] polyline arrays, then the envelope scratch
struct FPlan {
    int cap8;  // bytes of one double[cap] array (16-byte aligned)
    EnvScratch env;
    __host__ __device__ FPlan(int cap, int isec_cap) {
        cap8 = EnvScratch::align16(cap * 8);
        env.off = 14 * cap8;
        env.cap = cap;
        env.isec_cap = isec_cap;
    }
    __host__ __device__ size_t bytes() const { return (size_t)env.off + env.bytes(); }
};

size_t solver_f_smem_bytes(int cap, int isec_cap) { return FPlan(cap, isec_cap).bytes(); }


#ifdef DCEGM_PHASE_CLOCKS
__device__ unsigned long long dc_phase_total[kPhaseSlots];
#endif

__global__ void __launch_bounds__(kThreads, kCtasPerSm) solve_f_kernel(DeviceBatch B) {
    const int m = B.order ? B.order[B.first + blockIdx.x] : B.first + blockIdx.x;
    const int tid = threadIdx.x;
#ifdef DCEGM_PHASE_CLOCKS
    if (tid == 0) {
        for (int i = 0; i < kPhaseSlots; i++) dc_clk()[i] = 0;
        dc_clk()[kPhaseSlots] = clock64();
    }
#endif
    __shared__ DevModel s_p;
    if (tid == 0) s_p = B.models[m];
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const FPlan P(B.lay.cap, B.lay.isec_cap);
    const EnvScratch S = P.env;
    LineSet C, N;  // C is being built (period t), N holds period t+1
    C.cap8 = N.cap8 = P.cap8;
    C.base = 0;
    N.base = 7 * P.cap8;
    __shared__ double s_ev0[2];
    __shared__ int s_n[2];
    __shared__ int s_status;

    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny;
    const double* __restrict__ offm = B.off + (size_t)m * (nT - 1) * 2 * ny;
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const double a0 = avec[0];
    const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
    const bool can_skip = p.cfloor > 0.0;  // 0*up(c) == 0 exactly when c >= cfloor > 0
    // the fast EGM path needs finite exponents and a positive normal consumption floor
    const bool model_fast = isfinite(p.neg_gamma) && isfinite(p.oneminusgamma) && isfinite(p.neg_oneover_gamma) &&
                            isfinite(p.oneover_oneminusgamma) && dc_is_norm_pos(p.cfloor) && p.lambda > 1e-100 && p.lambda < 1e100;
    // knot-index tables alias the envelope's Tx/Ty arrays (free during the EGM step)
    const int tab0 = S.o_d(3);
#ifdef DC_KNOT_WIDE  // the tables may run over the whole envelope scratch behind the dirty line (free during the EGM step)
    const int nbuck = knot_index_buckets((int)S.bytes() - 3 * P.cap8, 3);
#else
    const int nbuck = knot_index_buckets(2 * P.cap8, 3);
#endif
    const int tab_stride = (nbuck + 8) * 2;
    // this period's income offsets and the quadrature weights, staged once (the EGM loops read them per node)
    __shared__ double s_off[2 * kMaxNy], s_wgt[kMaxNy];
    if (tid < ny) s_wgt[tid] = wgt[tid];
    // this thread's savings grid points stay in registers across the periods
    const double a_reg0 = tid < na ? avec[tid] : 0.0;
    const double a_reg1 = tid + kThreads < na ? avec[tid + kThreads] : 0.0;

    if (tid == 0) s_status = 0;

    // summary accumulators (thread 0)
    int sum_envp = 0, sum_seg = 0, sum_is = 0, sum_rm = 0, sum_maxlen = 0;
    long long sum_pts = 0;

    for (int t = nT - 1; t >= 0; t--) {
        if (t == nT - 1) {
            // terminal period (dc_algo.jl:14-20)
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
                // retiree terminal value line lives on the savings grid, not on r_x: stage through D
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
                sum_pts += 4 + 2 * nv;
            }
            __syncthreads();
            {
                const LineSet tmp = N;
                N = C;
                C = tmp;
            }
            continue;
        }

        DC_CLK(0);
        const bool next_terminal = (t + 1 == nT - 1);
        const bool beq = (t == nT - 2) && p.nu > 0;
        const double scale = beq ? p.betaRnu : p.betaR;
        if (tid < 2 * ny) s_off[tid] = offm[(size_t)t * 2 * ny + tid];  // visible after the barrier below

        // ---------------- 1. EGM step -----------------------------------------------------------------
        // fast path: straight-line node evaluation (bucket-indexed knot search, pow from shared
        // logarithms); see egm_node_generic for the conditions under which it does not apply
        bool fast = model_fast && !beq && !(s_status & kStatusNoIndex) && N.n_wc >= 2 &&
                          N.n_wv >= 2 && N.n_r >= 2;
        KnotIndex Kr, Kc, Kv;
        Kr.tab = tab0;
        Kc.tab = tab0 + tab_stride;
        Kv.tab = tab0 + 2 * tab_stride;
        Kr.lo = Kc.lo = Kv.lo = Kr.inv = Kc.inv = Kv.inv = 0.0;
        Kr.topi = Kc.topi = Kv.topi = 0;
        if (fast) {
            knot_index_range(Kr, N.r_x(), N.n_r, nbuck);
            knot_index_range(Kc, N.wc_x(), N.n_wc, nbuck);
            knot_index_range(Kv, N.wv_x(), N.n_wv, nbuck);
#ifdef DC_KNOT_BUILD_PER_KNOT  // A/B builds only
            const int nmax = max(N.n_r, max(N.n_wc, N.n_wv));
            for (int k = tid; k < nmax; k += kThreads) {  // the three tables in one pass
                if (k < N.n_r) knot_index_build_step(Kr, N.r_x(), N.n_r, nbuck, k);
                if (k < N.n_wc) knot_index_build_step(Kc, N.wc_x(), N.n_wc, nbuck, k);
                if (k < N.n_wv) knot_index_build_step(Kv, N.wv_x(), N.n_wv, nbuck, k);
            }
#else
            knot_index_build3<kThreads>(reinterpret_cast<unsigned short*>(dc_smem + Kr.tab),
                                        reinterpret_cast<unsigned short*>(dc_smem + Kc.tab),
                                        reinterpret_cast<unsigned short*>(dc_smem + Kv.tab), Kr, Kc, Kv, N.r_x(), N.n_r,
                                        N.wc_x(), N.n_wc, N.wv_x(), N.n_wv, nbuck);
#endif
        }
        __syncthreads();
        DC_CLK(1);
        bool same_nodes = true;  // retiree: do all nodes see the same next-period cash-on-hand?
        for (int iy = 1; iy < ny; iy++) same_nodes &= s_off[ny + iy] == s_off[ny];
        if (fast) {  // all queries of this period inside the range the bucket map handles exactly
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

        // ---- thread per savings grid point: worker today (id = 0), serial over the income nodes ----
        {
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
            const bool same = same_nodes && fast_r;  // one evaluation serves all nodes (same inputs, same bits)
            for (int ia = tid; ia < na; ia += kThreads) {
                const double a = ia == tid ? a_reg0 : (ia == tid + kThreads ? a_reg1 : avec[ia]);
#define EGM_COUNT_LE(K, x, n, q) count_le_indexed(K, x, n, q)
#define EGM_LINESET LineSet
#include "egm_point.inc"
#undef EGM_COUNT_LE
#undef EGM_LINESET
            }
        }
        DC_CLK(2);
        __syncthreads();
        DC_CLK(3);
        const double ev0w = s_ev0[0], ev0r = s_ev0[1];

        if (B.dirty) {  // runfdemo: vdirty / cdirty (dc_algo.jl:79-82)
            double2* d = B.dirty + line_slot(B.lay, m, t, 0, 0, 0) * (size_t)na;
            for (int k = tid; k < na; k += kThreads) {
                d[k] = make_double2(S.Dx()[koff + k], S.Dc()[koff + k]);                 // worker c
                d[(size_t)na + k] = make_double2(S.Dx()[koff + k], S.Dv()[koff + k]);    // worker v
                d[(size_t)2 * na + k] = make_double2(C.r_x()[1 + k], C.rc_y()[1 + k]);   // retiree c
                d[(size_t)3 * na + k] = make_double2(C.r_x()[1 + k], C.rv_y()[1 + k]);   // retiree v
            }
        }

        // ---------------- 2. worker: secondary envelope + policy repair -----------------------------
        const size_t ss = state_slot(B.lay, m, t, 0);
        // lines of the first period are never interpolated by the solve (dc_algo.jl:22-37 reads it+1 only)
        const bool first_period = t == 0;
        worker_envelope(S, koff, na, a0, ev0w, p, C.wv_x(), C.wv_y(), &s_n[0], C.wc_x(), C.wc_y(), &s_n[1], B.meta + ss,
                        B.isects + ss * (size_t)B.lay.isec_cap, B.removed + ss * (size_t)B.lay.rem_words,
                        B.lay.rem_words, first_period);
        C.n_wv = s_n[0];
        C.n_wc = s_n[1];
        C.vb_w = ev0w;
        DC_CLK(12);

        // ---------------- 3. retiree: credit-constraint point and checks (dc_algo.jl:90-101) --------
        if (tid == 0) {
            C.r_x()[0] = a0;
            C.rc_y()[0] = 0.0;
            C.rv_y()[0] = ev0r;
        }
        C.n_r = na + 1;
        C.vb_r = ev0r;
        __syncthreads();
        int flag = 0;
        for (int k = tid; k < na + 1; k += kThreads) {
            if (k + 1 < na + 1) flag |= knot_pair_flag(C.r_x()[k], C.r_x()[k + 1], first_period);
            if (!isfinite(C.r_x()[k]) || !isfinite(C.rc_y()[k]) || !isfinite(C.rv_y()[k])) flag |= DCEGM_ST_NONFINITE;
        }
        if (flag) atomicOr(&s_status, flag);
        if (tid == 0 && S.cnt()[2]) atomicOr(&s_status, S.cnt()[2]);

        store_line(B, m, t, 0, 0, DCEGM_KIND_C, C.wc_x(), C.wc_y(), C.n_wc);
        store_line(B, m, t, 0, 0, DCEGM_KIND_V, C.wv_x(), C.wv_y(), C.n_wv);
        store_line(B, m, t, 0, 1, DCEGM_KIND_C, C.r_x(), C.rc_y(), C.n_r);
        store_line(B, m, t, 0, 1, DCEGM_KIND_V, C.r_x(), C.rv_y(), C.n_r);
        if (tid == 0) {
            B.vbound[ss * 2 + 0] = ev0w;
            B.vbound[ss * 2 + 1] = ev0r;
            const dcegm_env_meta mt = B.meta[ss];
            sum_envp += mt.n_removed > 0;
            sum_seg += mt.n_segments;
            sum_is += mt.n_isects;
            sum_rm += mt.n_removed;
            int ml = C.n_wc > C.n_wv ? C.n_wc : C.n_wv;
            ml = ml > C.n_r ? ml : C.n_r;
            sum_maxlen = sum_maxlen > ml ? sum_maxlen : ml;
            sum_pts += C.n_wc + C.n_wv + 2 * C.n_r;
        }
        __syncthreads();
        DC_CLK(13);
        {
            const LineSet tmp = N;
            N = C;
            C = tmp;
        }
    }

    // ---------------- per-model summary (payload of the multi-GPU allgather) -------------------------
    const LineSet& L0 = N;  // period 1
    double cs_c = 0.0, cs_v = 0.0;
    for (int k = tid; k < L0.n_wc; k += kThreads) cs_c += L0.wc_y()[k];
    for (int k = tid; k < L0.n_r; k += kThreads) cs_c += L0.rc_y()[k];
    for (int k = tid; k < L0.n_wv; k += kThreads) cs_v += isfinite(L0.wv_y()[k]) ? L0.wv_y()[k] : 0.0;
    for (int k = tid; k < L0.n_r; k += kThreads) cs_v += isfinite(L0.rv_y()[k]) ? L0.rv_y()[k] : 0.0;
    // fixed-order block sums: warp butterfly then ascending warps
    __shared__ double s_cs[2][kWarps];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cs_c += __shfl_xor_sync(0xffffffffu, cs_c, o);
        cs_v += __shfl_xor_sync(0xffffffffu, cs_v, o);
    }
    if ((tid & 31) == 0) {
        s_cs[0][tid >> 5] = cs_c;
        s_cs[1][tid >> 5] = cs_v;
    }
    __syncthreads();
    if (tid == 0) {
        dcegm_summary sm;
        sm.status = s_status;
        sm.n_env_periods = sum_envp;
        sm.tot_segments = sum_seg;
        sm.tot_isects = sum_is;
        sm.tot_removed = sum_rm;
        sm.max_len = sum_maxlen;
        sm.tot_points = sum_pts;
        const double fr[DCEGM_SUMMARY_NPROBE] = {0.02, 0.05, 0.1, 0.2, 0.3, 0.5, 0.7, 0.9};
        for (int q = 0; q < DCEGM_SUMMARY_NPROBE; q++) {
            const double mq = p.a_low + fr[q] * (p.a_high - p.a_low);
            double c = interp_y(L0.wc_x(), L0.wc_y(), L0.n_wc, mq);
            c = c < p.cfloor ? p.cfloor : c;
            sm.c1_probe[q] = c;
            sm.v1_probe[q] = nT >= 2 ? vfun_dev(true, false, c, mq, L0.wv_x(), L0.wv_y(), L0.n_wv, L0.vb_w, p) : 0.0;
        }
        double a = 0.0, b = 0.0;
        for (int w = 0; w < kWarps; w++) {
            a += s_cs[0][w];
            b += s_cs[1][w];
        }
        sm.checksum_c = a;
        sm.checksum_v = b;
        B.summaries[m] = sm;
        B.status[m] = s_status;
    }
    signal_model_done(B, m);
#ifdef DCEGM_PHASE_CLOCKS
    DC_CLK(14);
    if (tid == 0)
        for (int i = 0; i < kPhaseSlots; i++) atomicAdd(&dc_phase_total[i], (unsigned long long)dc_clk()[i]);
#endif
}

#ifdef DCEGM_PHASE_CLOCKS
// profiling builds only: cycles per phase summed over all CTAs since the last reset
extern "C" int dcegm_debug_phase_clocks(unsigned long long* out, int reset) {
    if (out && cudaMemcpyFromSymbol(out, dc_phase_total, sizeof(unsigned long long) * kPhaseSlots) != cudaSuccess) return -2;
    if (reset) {
        unsigned long long z[kPhaseSlots] = {0};
        if (cudaMemcpyToSymbol(dc_phase_total, z, sizeof z) != cudaSuccess) return -2;
    }
    return 0;
}
#endif

cudaError_t launch_solve_f(const DeviceBatch& B0, int first, int count, cudaStream_t stream) {
    DeviceBatch B = B0;
    B.first = first;
    const size_t smem = FPlan(B.lay.cap, B.lay.isec_cap).bytes();
    // the attribute is per device and the call is cheap: set on every launch (several handles / GPUs per process)
    cudaError_t e = cudaFuncSetAttribute(solve_f_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    solve_f_kernel<<<count, kThreads, smem, stream>>>(B);
    return cudaGetLastError();
}

}  // namespace dcegm
