// solver_g.cu — persistent per-model kernel for the general model with state-dependent income.
//
// Replaces dc_EGM!(m::GModel, p) (src/dc_algo.jl:112-206).  One CTA owns one parameterisation.
// The 2 x ny next-period line pairs (policy/value x worker/retiree for every income state) stay in
// shared memory; once the expectations of a period are formed they are no longer needed, so the
// finished lines of the current period overwrite them in place — one resident set, every line
// written to HBM once.  Per period:
//   A. thread per (current state iy, future choice iid, grid point ia): loop over the future state
//      jy, accumulate pr*vfun and pr*u'(c') in index order (dc_algo.jl:142-157)
//   B. thread per (current choice id, ia) of one state: ccp / logsum on the integrated values, Euler
//      inversion, value (dc_algo.jl:159-182)
//   C. worker envelope + policy repair (envelope.cuh), credit-constraint point, store.
// B and C run per income state on THREAD GROUPS (G = 2 or 4 groups of 256 / G threads, each with its own envelope
// scratch and named barrier): the envelopes of a period — short lines, many barriers, stretches where one warp works —
// proceed side by side instead of one after the other on the whole CTA.
#include <math_constants.h>

#include <cstdlib>

#define DC_SEG_INTERP_ATTR static __device__ __noinline__  // envelope.cuh
#define DC_GROUPED 1  // block_* and worker_envelope run on groups of NT threads (device_common.cuh)
#include "device_common.cuh"
#include "envelope.cuh"

namespace dcegm {

// shared-memory map: ny x 4 double[cap] worker line arrays | ny x 3 double[na + 1] retiree line arrays |
// vmat,cmat double[ny][2][na] each | per-state words | `groups` envelope scratch regions (the dirty worker line of a
// state is written straight into its group's region)
struct GPlan {
    int cap8, na8, r8, ny, groups;
    int o_ret, o_mat, o_meta, env_stride;
    EnvScratch env;  // region of group 0
    __host__ __device__ GPlan(int cap, int isec_cap, int na, int ny_, int groups_) {
        ny = ny_;
        groups = groups_;
        cap8 = EnvScratch::align16(cap * 8);
        na8 = EnvScratch::align16(na * 8);
        r8 = EnvScratch::align16((na + 1) * 8);  // the retiree's lines have exactly na + 1 points
        o_ret = ny * 4 * cap8;
        o_mat = o_ret + ny * 3 * r8;
        o_meta = o_mat + ny * 4 * na8;
        env.off = o_meta + EnvScratch::align16(ny * 5 * 8) + EnvScratch::align16(ny * 3 * 4);
        env.cap = cap;
        env.isec_cap = isec_cap;
        env_stride = EnvScratch::align16(env.bytes());
    }
    __host__ __device__ size_t bytes() const { return (size_t)env.off + (size_t)groups * env_stride; }
    __device__ __forceinline__ EnvScratch scratch(int g) const {
        EnvScratch S = env;
        S.off += g * env_stride;
        return S;
    }
    // line arrays of income state iy: 0 wc_x, 1 wc_y, 2 wv_x, 3 wv_y, 4 r_x, 5 rc_y, 6 rv_y
    __device__ __forceinline__ double* line(int iy, int a) const {
        return reinterpret_cast<double*>(dc_smem + (a < 4 ? (iy * 4 + a) * cap8 : o_ret + (iy * 3 + a - 4) * r8));
    }
    // 0 vmat[iid=0], 1 vmat[iid=1], 2 cmat[0], 3 cmat[1]
    __device__ __forceinline__ double* mat(int iy, int a) const {
        return reinterpret_cast<double*>(dc_smem + o_mat + (iy * 4 + a) * na8);
    }
    // per state: vbound worker, vbound retiree, ev0 worker, ev0 retiree (doubles); n_wc, n_wv, n_r (ints)
    __device__ __forceinline__ double* vb(int iy) const { return reinterpret_cast<double*>(dc_smem + o_meta) + iy * 5; }
    __device__ __forceinline__ int* len(int iy) const {
        return reinterpret_cast<int*>(dc_smem + o_meta + EnvScratch::align16(ny * 5 * 8)) + iy * 3;
    }
};

// what the kernel needs at least (two thread groups); launch_solve_g takes four when they fit twice per SM
size_t solver_g_smem_bytes(int cap, int isec_cap, int na, int ny) { return GPlan(cap, isec_cap, na, ny, 2).bytes(); }

// stored by the nt threads t = 0 .. nt - 1 of the CTA or of one thread group
__device__ inline void store_line_g(const DeviceBatch& B, int m, int it, int iy, int id, int kind, const double* x,
                                    const double* y, int n, int t = threadIdx.x, int nt = kThreads) {
    const size_t slot = line_slot(B.lay, m, it, iy, id, kind);
    if (t == 0) B.lens[slot] = n;
    if (!B.lines) return;
    double2* dst = B.lines + slot * (size_t)B.lay.cap;
    const int nn = n < B.lay.cap ? n : B.lay.cap;
    for (int k = t; k < nn; k += nt) dst[k] = make_double2(x[k], y[k]);
}

// Top = largest step of the fixed-trip knot search (lines hold fewer than 2*Top points; chosen at launch from cap)
// G = thread groups (income states in flight through phases B and C), kThreads / G threads each
template <int Top, int G>
__global__ void __launch_bounds__(kThreads, kCtasPerSm) solve_g_kernel(DeviceBatch B) {
    constexpr int kGroups = G, kGroupThreads = kThreads / G;
    static_assert(kGroupThreads % 32 == 0 && kGroups * kGroupThreads == kThreads && kGroups <= 8, "thread groups");
    const int m = B.order ? B.order[B.first + blockIdx.x] : B.first + blockIdx.x;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    __shared__ int s_status;
    __shared__ int s_n[kGroups][2];
    __shared__ int s_sum[6];
    const int grp = tid / kGroupThreads, gt = tid % kGroupThreads;  // thread group of phases B and C, index inside it
    if (tid == 0) {
        s_p = B.models[m];
        s_status = 0;
    }
    if (tid < 6) s_sum[tid] = 0;
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const GPlan P(B.lay.cap, B.lay.isec_cap, na, ny, kGroups);
    const EnvScratch S = P.scratch(grp);  // this thread group's envelope scratch
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny * ny;               // [iy][jy]
    const double* __restrict__ offm = B.off + (size_t)m * (nT - 1) * 2 * ny;    // [t][iid][jy]
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const double a0 = avec[0];
    const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
    int sum_envp = 0, sum_seg = 0, sum_is = 0, sum_rm = 0, sum_maxlen = 0;
    long long sum_pts = 0;

    // ---- terminal period (dc_algo.jl:122-131) ------------------------------------------------------
    {
        const int t = nT - 1;
        const int nv = p.nu > 0 ? na : 2;
        const EnvScratch S0 = P.scratch(0);  // staging of the retiree's bequest line (whole CTA)
        for (int iy = 0; iy < ny; iy++) {
            if (tid == 0) {
                double* wcx = P.line(iy, 0); double* wcy = P.line(iy, 1);
                double* rx = P.line(iy, 4);  double* rcy = P.line(iy, 5);
                wcx[0] = p.a_lowT; wcy[0] = 0.0; wcx[1] = p.a_high; wcy[1] = p.a_high;
                rx[0] = p.a_lowT;  rcy[0] = 0.0; rx[1] = p.a_high;  rcy[1] = p.a_high;
                if (!(p.nu > 0)) {
                    double* wvx = P.line(iy, 2); double* wvy = P.line(iy, 3); double* rvy = P.line(iy, 6);
                    wvx[0] = p.a_lowT; wvy[0] = 0.0; wvx[1] = p.a_high; wvy[1] = kNaN;
                    rvy[0] = 0.0; rvy[1] = kNaN;
                }
                P.len(iy)[0] = 2; P.len(iy)[1] = nv; P.len(iy)[2] = 2;
                P.vb(iy)[0] = 0.0; P.vb(iy)[1] = 0.0;
            }
            if (p.nu > 0) {
                // m.v[id,iy,nT] = bequest(avec, id==1) - 10*(id>1) on the savings grid; never read by vfun
                for (int k = tid; k < na; k += kThreads) {
                    const double b = p.nu * u_plain(avec[k] + p.bbar, p);
                    P.line(iy, 2)[k] = avec[k];
                    P.line(iy, 3)[k] = (b - p.alpha) - 0.0;
                    S0.Dx()[k] = avec[k];
                    S0.Dv()[k] = (b - 0.0) - 10.0;
                }
            }
            __syncthreads();
            store_line_g(B, m, t, iy, 0, DCEGM_KIND_C, P.line(iy, 0), P.line(iy, 1), 2);
            store_line_g(B, m, t, iy, 1, DCEGM_KIND_C, P.line(iy, 4), P.line(iy, 5), 2);
            store_line_g(B, m, t, iy, 0, DCEGM_KIND_V, P.line(iy, 2), P.line(iy, 3), nv);
            if (p.nu > 0)
                store_line_g(B, m, t, iy, 1, DCEGM_KIND_V, S0.Dx(), S0.Dv(), na);
            else
                store_line_g(B, m, t, iy, 1, DCEGM_KIND_V, P.line(iy, 4), P.line(iy, 6), 2);
            if (tid == 0) {
                const size_t ss = state_slot(B.lay, m, t, iy);
                B.vbound[ss * 2 + 0] = 0.0;
                B.vbound[ss * 2 + 1] = 0.0;
                dcegm_env_meta z = {0, 0, 0, 0};
                B.meta[ss] = z;
                sum_pts += 4 + 2 * nv;
            }
            __syncthreads();
        }
    }

    for (int t = nT - 2; t >= 0; t--) {
        const bool next_terminal = (t + 1 == nT - 1);
        const bool beq = (t == nT - 2) && p.nu > 0;  // is_bequest, dc_algo.jl:120
        const double scale = beq ? p.betaRnu : p.betaR;
        const double* __restrict__ off_t = offm + (size_t)t * 2 * ny;

        // ---- A. expectations over the future state (dc_algo.jl:142-157) ----------------------------
        for (int item = tid; item < ny * 2 * na; item += kThreads) {
            const int ia = item % na;
            const int iid = (item / na) & 1;
            const int iy = item / (2 * na);
            const double aR = avec[ia] * p.R;
            double vacc = 0.0, cacc = 0.0;  // fill!(vmat,0.0); fill!(cmat,0.0)
            for (int jy = 0; jy < ny; jy++) {
                const double pr = wgt[iy * ny + jy];
                const double m1 = aR + off_t[iid * ny + jy];  // param.jl:301
                double c1, v1;
                if (iid == 0) {
                    const double* cx = P.line(jy, 0); const double* cy = P.line(jy, 1);
                    const int ncw = P.len(jy)[0];
                    c1 = interp_y<Top>(cx, cy, ncw, m1);
                    c1 = c1 < p.cfloor ? p.cfloor : c1;
                    if (next_terminal && p.nu > 0)
                        v1 = p.nu * u_plain(c1 + p.bbar, p) - 0.0;
                    else if (next_terminal || m1 < P.line(jy, 2)[1])
                        v1 = u_work(c1, true, p) + p.beta * P.vb(jy)[0];
                    else
                        v1 = interp_y<Top>(P.line(jy, 2), P.line(jy, 3), P.len(jy)[1], m1);
                } else {
                    const double* rx = P.line(jy, 4);
                    const int nr = P.len(jy)[2];
                    const int ir = left_knot<Top>(rx, nr, m1);
                    c1 = interp_at(rx, P.line(jy, 5), nr, ir, m1);
                    c1 = c1 < p.cfloor ? p.cfloor : c1;
                    if (next_terminal && p.nu > 0)
                        v1 = p.nu * u_plain(c1 + p.bbar, p) - 0.0;
                    else if (next_terminal || m1 < rx[1])
                        v1 = u_work(c1, false, p) + p.beta * P.vb(jy)[1];
                    else
                        v1 = interp_at(rx, P.line(jy, 6), nr, ir, m1);
                }
                vacc = vacc + pr * v1;                                   // dc_algo.jl:149
                cacc = cacc + pr * up(beq ? c1 + p.bbar : c1, p);       // dc_algo.jl:150-154
            }
            P.mat(iy, iid)[ia] = vacc;
            P.mat(iy, 2 + iid)[ia] = cacc;
        }
        __syncthreads();

        // ---- B + C, one income state per thread group -------------------------------------------------
        for (int iy0 = 0; iy0 < ny; iy0 += kGroups) {
            const int iy = iy0 + grp;
            if (iy >= ny) continue;  // uniform over the group: its named barrier is not used in this round
            // ---- B. current choice: ccp / logsum on integrated values, Euler inversion (dc_algo.jl:159-182)
            for (int item = gt; item < 2 * na; item += kGroupThreads) {
                const int ia = item % na;
                const int id = item / na;
                const double v1 = P.mat(iy, 0)[ia], v2 = P.mat(iy, 1)[ia];
                const double cm1 = P.mat(iy, 2)[ia], cm2 = P.mat(iy, 3)[ia];
                double pw = 0.0, ev = v2;
                if (id == 0) {
                    const bool w_hi = v1 > v2;
                    const double mx = w_hi ? v1 : v2;
                    const double d = ((w_hi ? v2 : v1) - mx) / p.lambda;
                    const double e = dc_exp(d);
                    const double e1 = w_hi ? 1.0 : e, e2 = w_hi ? e : 1.0;
                    const double s = e1 + e2;
                    pw = e1 / s;
                    ev = mx + p.lambda * dc_log(s);
                }
                const double mu1 = pw * cm1 + (1.0 - pw) * cm2;  // dc_algo.jl:165
                const double rhs = scale * mu1;
                const double c0 = iup(rhs, p);
                const double x = avec[ia] + c0;
                const double vy = u_work(c0, id == 0, p) + p.beta * ev;
                if (id == 0) {  // the dirty worker line, where worker_envelope expects it
                    S.Dx()[koff + ia] = x;
                    S.Dv()[koff + ia] = vy;
                    S.Dc()[koff + ia] = c0;
                    if (ia == 0) P.vb(iy)[2] = ev;
                } else {
                    // the t+1 lines are dead after phase A: the retiree lines of t are built in place
                    P.line(iy, 4)[1 + ia] = x;
                    P.line(iy, 5)[1 + ia] = c0;
                    P.line(iy, 6)[1 + ia] = vy;
                    if (ia == 0) P.vb(iy)[3] = ev;
                }
            }
            dc_gsync<kGroupThreads>();

            // ---- C. envelope, credit-constraint point, checks, store ------------------------------------
            const size_t ss = state_slot(B.lay, m, t, iy);
            const bool first_period = t == 0;  // see solve_f_kernel
            const double ev0w = P.vb(iy)[2], ev0r = P.vb(iy)[3];
            if (B.dirty) {
                double2* d = B.dirty + line_slot(B.lay, m, t, iy, 0, 0) * (size_t)na;
                for (int k = gt; k < na; k += kGroupThreads) {
                    d[k] = make_double2(S.Dx()[koff + k], S.Dc()[koff + k]);
                    d[(size_t)na + k] = make_double2(S.Dx()[koff + k], S.Dv()[koff + k]);
                    d[(size_t)2 * na + k] = make_double2(P.line(iy, 4)[1 + k], P.line(iy, 5)[1 + k]);
                    d[(size_t)3 * na + k] = make_double2(P.line(iy, 4)[1 + k], P.line(iy, 6)[1 + k]);
                }
                dc_gsync<kGroupThreads>();  // read before the envelope prepends bend-back points in front
            }
            worker_envelope<EnvScratch, kGroupThreads>(S, koff, na, a0, ev0w, p, P.line(iy, 2), P.line(iy, 3), &s_n[grp][0],
                                                       P.line(iy, 0), P.line(iy, 1), &s_n[grp][1], B.meta + ss,
                                                       B.isects + ss * (size_t)B.lay.isec_cap,
                                                       B.removed + ss * (size_t)B.lay.rem_words, B.lay.rem_words, first_period);
            const int n_wv = s_n[grp][0], n_wc = s_n[grp][1], n_r = na + 1;
            if (gt == 0) {
                P.line(iy, 4)[0] = a0;
                P.line(iy, 5)[0] = 0.0;
                P.line(iy, 6)[0] = ev0r;
                P.len(iy)[0] = n_wc;
                P.len(iy)[1] = n_wv;
                P.len(iy)[2] = n_r;
                P.vb(iy)[0] = ev0w;
                P.vb(iy)[1] = ev0r;
            }
            dc_gsync<kGroupThreads>();
            int flag = 0;
            {
                const double* rx = P.line(iy, 4);
                for (int k = gt; k < n_r; k += kGroupThreads) {
                    if (k + 1 < n_r) flag |= knot_pair_flag(rx[k], rx[k + 1], first_period);
                    if (!isfinite(rx[k]) || !isfinite(P.line(iy, 5)[k]) || !isfinite(P.line(iy, 6)[k])) flag |= DCEGM_ST_NONFINITE;
                }
            }
            if (flag) atomicOr(&s_status, flag);
            if (gt == 0 && S.cnt()[2]) atomicOr(&s_status, S.cnt()[2]);
            store_line_g(B, m, t, iy, 0, DCEGM_KIND_C, P.line(iy, 0), P.line(iy, 1), n_wc, gt, kGroupThreads);
            store_line_g(B, m, t, iy, 0, DCEGM_KIND_V, P.line(iy, 2), P.line(iy, 3), n_wv, gt, kGroupThreads);
            store_line_g(B, m, t, iy, 1, DCEGM_KIND_C, P.line(iy, 4), P.line(iy, 5), n_r, gt, kGroupThreads);
            store_line_g(B, m, t, iy, 1, DCEGM_KIND_V, P.line(iy, 4), P.line(iy, 6), n_r, gt, kGroupThreads);
            if (gt == 0) {
                B.vbound[ss * 2 + 0] = ev0w;
                B.vbound[ss * 2 + 1] = ev0r;
                const dcegm_env_meta mt = B.meta[ss];
                sum_envp += mt.n_removed > 0;
                sum_seg += mt.n_segments;
                sum_is += mt.n_isects;
                sum_rm += mt.n_removed;
                int ml = n_wc > n_wv ? n_wc : n_wv;
                ml = ml > n_r ? ml : n_r;
                sum_maxlen = sum_maxlen > ml ? sum_maxlen : ml;
                sum_pts += n_wc + n_wv + 2 * n_r;
            }
            dc_gsync<kGroupThreads>();
        }
        __syncthreads();  // every state's lines of period t are in place before phase A of t - 1 reads them
    }

    // the group leaders' partial sums -> thread 0
    if (gt == 0 && grp > 0) {
        atomicAdd(&s_sum[0], sum_envp);
        atomicAdd(&s_sum[1], sum_seg);
        atomicAdd(&s_sum[2], sum_is);
        atomicAdd(&s_sum[3], sum_rm);
        atomicMax(&s_sum[4], sum_maxlen);
        atomicAdd(&s_sum[5], (int)sum_pts);
    }
    __syncthreads();
    if (tid == 0) {
        sum_envp += s_sum[0];
        sum_seg += s_sum[1];
        sum_is += s_sum[2];
        sum_rm += s_sum[3];
        sum_maxlen = sum_maxlen > s_sum[4] ? sum_maxlen : s_sum[4];
        sum_pts += s_sum[5];
    }

    // ---- per-model summary ----------------------------------------------------------------------------
    if (tid == 0) {
        dcegm_summary sm;
        memset(&sm, 0, sizeof sm);
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
            double c = interp_y(P.line(0, 0), P.line(0, 1), P.len(0)[0], mq);
            c = c < p.cfloor ? p.cfloor : c;
            sm.c1_probe[q] = c;
            sm.v1_probe[q] = (mq < P.line(0, 2)[1]) ? u_work(c, true, p) + p.beta * P.vb(0)[0]
                                                   : interp_y(P.line(0, 2), P.line(0, 3), P.len(0)[1], mq);
        }
        double cs_c = 0.0, cs_v = 0.0;
        for (int iy = 0; iy < ny; iy++) {
            for (int k = 0; k < P.len(iy)[0]; k++) cs_c += P.line(iy, 1)[k];
            for (int k = 0; k < P.len(iy)[2]; k++) cs_c += P.line(iy, 5)[k];
            for (int k = 0; k < P.len(iy)[1]; k++) cs_v += isfinite(P.line(iy, 3)[k]) ? P.line(iy, 3)[k] : 0.0;
            for (int k = 0; k < P.len(iy)[2]; k++) cs_v += isfinite(P.line(iy, 6)[k]) ? P.line(iy, 6)[k] : 0.0;
        }
        sm.checksum_c = cs_c;
        sm.checksum_v = cs_v;
        B.summaries[m] = sm;
        B.status[m] = s_status;
    }
    signal_model_done(B, m);
}

template <int Top, int G>
static cudaError_t launch_solve_g_as(const DeviceBatch& B, int count, size_t smem, cudaStream_t stream) {
    cudaError_t e = cudaFuncSetAttribute(solve_g_kernel<Top, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    solve_g_kernel<Top, G><<<count, kThreads, smem, stream>>>(B);
    return cudaGetLastError();
}

cudaError_t launch_solve_g(const DeviceBatch& B0, int first, int count, cudaStream_t stream) {
    DeviceBatch B = B0;
    B.first = first;
    // four thread groups when their scratch regions still leave room for kCtasPerSm CTAs per SM, else two
    const size_t smem4 = GPlan(B.lay.cap, B.lay.isec_cap, B.lay.na, B.lay.ny, 4).bytes();
    const size_t smem2 = GPlan(B.lay.cap, B.lay.isec_cap, B.lay.na, B.lay.ny, 2).bytes();
    const bool small = B.lay.cap < 256;  // na <= ~170: eight search steps instead of eleven
    bool four = false;
    if (B.lay.ny > 2) {
        int nb = 0;
        cudaError_t e = small ? cudaFuncSetAttribute(solve_g_kernel<128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4)
                              : cudaFuncSetAttribute(solve_g_kernel<kSearchTop, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
        if (e == cudaSuccess)
            e = small ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, solve_g_kernel<128, 4>, kThreads, smem4)
                      : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, solve_g_kernel<kSearchTop, 4>, kThreads, smem4);
        if (e != cudaSuccess) cudaGetLastError();  // too large for one CTA even: two groups
        four = e == cudaSuccess && nb >= kCtasPerSm;
    }
    const char* env = getenv("DCEGM_G_GROUPS");  // 2 or 4: overrides the choice (measurements)
    if (env) four = atoi(env) == 4;
    if (four) return small ? launch_solve_g_as<128, 4>(B, count, smem4, stream) : launch_solve_g_as<kSearchTop, 4>(B, count, smem4, stream);
    return small ? launch_solve_g_as<128, 2>(B, count, smem2, stream) : launch_solve_g_as<kSearchTop, 2>(B, count, smem2, stream);
}

}  // namespace dcegm
