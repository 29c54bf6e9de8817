// solver_g_big.cu — fine-grid solver of the general model (state-dependent income): grids whose 2 x ny
// next-period line pairs do not fit into one SM's shared memory.
//
// Same algorithm, arithmetic and output buffers as solve_g_kernel (solver_g.cu) — dc_EGM!(m::GModel, p),
// src/dc_algo.jl:112-206 — with the work of one period spread over the GPU, like solver_f_big.cu does for the
// retirement model:
//   bigg_terminal_kernel  one CTA per model: terminal lines of every income state (dc_algo.jl:122-131)
//   bigg_egm_kernel       CTA per (model, income state iy, 256 savings grid points): expectations over the future
//                         state jy in index order (dc_algo.jl:142-157), then ccp / logsum on the integrated values,
//                         Euler inversion and today's values for both choices (dc_algo.jl:159-182)
//   bigg_env_kernel       one CTA per (model, iy): secondary upper envelope + policy repair (envelope.cuh on
//                         global-memory scratch), credit-constraint point, checks, the one store of the lines
//   bigg_summary_kernel   one CTA per model: the summary record (same summation order as solve_g_kernel)
// The working lines (SoA, two period sets, all states) and one envelope scratch per state live in a per-model
// region of global memory.  Results are bit-identical to solve_g_kernel (tests/test_gpu_parity_g_big.py).
#include <math_constants.h>

#include "device_common.cuh"
#include "egm_node.cuh"
#include "envelope.cuh"

namespace dcegm {

constexpr int kEnvThreadsG = 1024;
constexpr int kTopG = MemGlobal::kTop;

// per-model state that survives between the launches of one solve
struct BigGState {
    int n_wc[2][kMaxNy], n_wv[2][kMaxNy], n_r[2][kMaxNy];  // [line set][income state]
    double vb_w[2][kMaxNy], vb_r[2][kMaxNy];
    double ev0w[kMaxNy], ev0r[kMaxNy];
    int status;
    int sum_envp, sum_seg, sum_is, sum_rm, sum_maxlen;
    unsigned long long sum_pts;
};

struct BigGPlan {
    int cap8, ny;
    EnvScratchG env;    // geometry of one state's envelope scratch (off = 0 inside its own region)
    size_t env_bytes;   // bytes of one state's envelope region
    size_t o_env, o_state, bytes;
    __host__ __device__ BigGPlan(int cap, int isec_cap, int ny_) {
        ny = ny_;
        cap8 = EnvScratchG::align16(cap * 8);
        env.g = nullptr;
        env.off = 0;
        env.cap = cap;
        env.isec_cap = isec_cap;
        env_bytes = ((size_t)env.bytes() + 255) / 256 * 256;
        o_env = (size_t)2 * ny * 7 * cap8;
        o_state = o_env + (size_t)ny * env_bytes;
        bytes = o_state + ((sizeof(BigGState) + 255) / 256 * 256);
    }
    // line arrays of income state iy in set `which`: 0 wc_x, 1 wc_y, 2 wv_x, 3 wv_y, 4 r_x, 5 rc_y, 6 rv_y
    __device__ __forceinline__ double* line(unsigned char* g, int which, int iy, int a) const {
        return reinterpret_cast<double*>(g + ((size_t)(which * ny + iy) * 7 + a) * cap8);
    }
    __device__ __forceinline__ EnvScratchG scratch(unsigned char* g, int iy) const {
        EnvScratchG S = env;
        S.g = g + o_env + (size_t)iy * env_bytes;
        return S;
    }
    __device__ __forceinline__ BigGState* state(unsigned char* g) const { return reinterpret_cast<BigGState*>(g + o_state); }
};

size_t solver_g_big_scratch_bytes(int cap, int isec_cap, int ny) { return BigGPlan(cap, isec_cap, ny).bytes; }

template <int NT>
__device__ inline void store_line_big(const DeviceBatch& B, int m, int it, int iy, int id, int kind, const double* x,
                                      const double* y, int n) {
    const size_t slot = line_slot(B.lay, m, it, iy, id, kind);
    if (threadIdx.x == 0) B.lens[slot] = n;
    if (!B.lines) return;
    double2* dst = B.lines + slot * (size_t)B.lay.cap;
    const int nn = n < B.lay.cap ? n : B.lay.cap;
    for (int k = threadIdx.x; k < nn; k += NT) dst[k] = make_double2(x[k], y[k]);
}

// ---- terminal period (dc_algo.jl:122-131) -------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) bigg_terminal_kernel(DeviceBatch B, unsigned char* scratch, size_t stride) {
    const int m = B.first + blockIdx.x;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    if (tid == 0) s_p = B.models[m];
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const BigGPlan P(B.lay.cap, B.lay.isec_cap, ny);
    unsigned char* g = scratch + (size_t)blockIdx.x * stride;
    BigGState* st = P.state(g);
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
    const int t = nT - 1, w = t & 1;
    const int nv = p.nu > 0 ? na : 2;
    for (int iy = 0; iy < ny; iy++) {
        EnvScratchG S = P.scratch(g, iy);
        if (tid == 0) {
            double* wcx = P.line(g, w, iy, 0); double* wcy = P.line(g, w, iy, 1);
            double* rx = P.line(g, w, iy, 4);  double* rcy = P.line(g, w, iy, 5);
            wcx[0] = p.a_lowT; wcy[0] = 0.0; wcx[1] = p.a_high; wcy[1] = p.a_high;
            rx[0] = p.a_lowT;  rcy[0] = 0.0; rx[1] = p.a_high;  rcy[1] = p.a_high;
            if (!(p.nu > 0)) {
                double* wvx = P.line(g, w, iy, 2); double* wvy = P.line(g, w, iy, 3); double* rvy = P.line(g, w, iy, 6);
                wvx[0] = p.a_lowT; wvy[0] = 0.0; wvx[1] = p.a_high; wvy[1] = kNaN;
                rvy[0] = 0.0; rvy[1] = kNaN;
            }
            st->n_wc[w][iy] = 2; st->n_wv[w][iy] = nv; st->n_r[w][iy] = 2;
            st->vb_w[w][iy] = 0.0; st->vb_r[w][iy] = 0.0;
        }
        if (p.nu > 0) {
            // m.v[id,iy,nT] = bequest(avec, id==1) - 10*(id>1) on the savings grid; never read by vfun
            for (int k = tid; k < na; k += kThreads) {
                const double b = p.nu * u_plain(avec[k] + p.bbar, p);
                P.line(g, w, iy, 2)[k] = avec[k];
                P.line(g, w, iy, 3)[k] = (b - p.alpha) - 0.0;
                S.Dx()[k] = avec[k];
                S.Dv()[k] = (b - 0.0) - 10.0;
            }
        }
        __syncthreads();
        store_line_big<kThreads>(B, m, t, iy, 0, DCEGM_KIND_C, P.line(g, w, iy, 0), P.line(g, w, iy, 1), 2);
        store_line_big<kThreads>(B, m, t, iy, 1, DCEGM_KIND_C, P.line(g, w, iy, 4), P.line(g, w, iy, 5), 2);
        store_line_big<kThreads>(B, m, t, iy, 0, DCEGM_KIND_V, P.line(g, w, iy, 2), P.line(g, w, iy, 3), nv);
        if (p.nu > 0)
            store_line_big<kThreads>(B, m, t, iy, 1, DCEGM_KIND_V, S.Dx(), S.Dv(), na);
        else
            store_line_big<kThreads>(B, m, t, iy, 1, DCEGM_KIND_V, P.line(g, w, iy, 4), P.line(g, w, iy, 6), 2);
        if (tid == 0) {
            const size_t ss = state_slot(B.lay, m, t, iy);
            B.vbound[ss * 2 + 0] = 0.0;
            B.vbound[ss * 2 + 1] = 0.0;
            dcegm_env_meta z = {0, 0, 0, 0};
            B.meta[ss] = z;
        }
        __syncthreads();
    }
    if (tid == 0) {
        st->status = 0;
        st->sum_envp = st->sum_seg = st->sum_is = st->sum_rm = st->sum_maxlen = 0;
        st->sum_pts = (unsigned long long)ny * (4 + 2 * nv);
    }
}

// ---- EGM step of period t for one income state: thread per savings grid point ---------------------------
__global__ void __launch_bounds__(kThreads) bigg_egm_kernel(DeviceBatch B, unsigned char* scratch, size_t stride, int t) {
    const int lm = blockIdx.z, iy = blockIdx.y;
    const int m = B.first + lm;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    if (tid == 0) s_p = B.models[m];
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const BigGPlan P(B.lay.cap, B.lay.isec_cap, ny);
    unsigned char* g = scratch + (size_t)lm * stride;
    BigGState* st = P.state(g);
    EnvScratchG S = P.scratch(g, iy);
    const int wn = (t + 1) & 1, wc = t & 1;  // next-period set (read), this period's set (written)
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny * ny;               // [iy][jy]
    const double* __restrict__ off_t = B.off + (size_t)m * (nT - 1) * 2 * ny + (size_t)t * 2 * ny;  // [iid][jy]
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const bool next_terminal = (t + 1 == nT - 1);
    const bool beq = (t == nT - 2) && p.nu > 0;  // is_bequest, dc_algo.jl:120
    const double scale = beq ? p.betaRnu : p.betaR;
    const int ia = blockIdx.x * kThreads + tid;
    if (ia >= na) return;
    const double aR = avec[ia] * p.R;
    // ---- A. expectations over the future state (dc_algo.jl:142-157) ----
    double vm[2], cm[2];
#pragma unroll
    for (int iid = 0; iid < 2; iid++) {
        double vacc = 0.0, cacc = 0.0;  // fill!(vmat,0.0); fill!(cmat,0.0)
        for (int jy = 0; jy < ny; jy++) {
            const double pr = wgt[iy * ny + jy];
            const double m1 = aR + off_t[iid * ny + jy];  // param.jl:301
            double c1, v1;
            if (iid == 0) {
                const double* cx = P.line(g, wn, jy, 0);
                c1 = interp_y<kTopG>(cx, P.line(g, wn, jy, 1), st->n_wc[wn][jy], m1);
                c1 = c1 < p.cfloor ? p.cfloor : c1;
                if (next_terminal && p.nu > 0)
                    v1 = p.nu * u_plain(c1 + p.bbar, p) - 0.0;
                else if (next_terminal || m1 < P.line(g, wn, jy, 2)[1])
                    v1 = u_work(c1, true, p) + p.beta * st->vb_w[wn][jy];
                else
                    v1 = interp_y<kTopG>(P.line(g, wn, jy, 2), P.line(g, wn, jy, 3), st->n_wv[wn][jy], m1);
            } else {
                const double* rx = P.line(g, wn, jy, 4);
                const int nr = st->n_r[wn][jy];
                const int ir = left_knot<kTopG>(rx, nr, m1);
                c1 = interp_at(rx, P.line(g, wn, jy, 5), nr, ir, m1);
                c1 = c1 < p.cfloor ? p.cfloor : c1;
                if (next_terminal && p.nu > 0)
                    v1 = p.nu * u_plain(c1 + p.bbar, p) - 0.0;
                else if (next_terminal || m1 < rx[1])
                    v1 = u_work(c1, false, p) + p.beta * st->vb_r[wn][jy];
                else
                    v1 = interp_at(rx, P.line(g, wn, jy, 6), nr, ir, m1);
            }
            vacc = vacc + pr * v1;                               // dc_algo.jl:149
            cacc = cacc + pr * up(beq ? c1 + p.bbar : c1, p);   // dc_algo.jl:150-154
        }
        vm[iid] = vacc;
        cm[iid] = cacc;
    }
    // ---- B. current choice: ccp / logsum on the integrated values, Euler inversion (dc_algo.jl:159-182) ----
#pragma unroll
    for (int id = 0; id < 2; id++) {
        const double v1 = vm[0], v2 = vm[1];
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
        const double mu1 = pw * cm[0] + (1.0 - pw) * cm[1];  // dc_algo.jl:165
        const double rhs = scale * mu1;
        const double c0 = iup(rhs, p);
        const double x = avec[ia] + c0;
        const double vy = u_work(c0, id == 0, p) + p.beta * ev;
        if (id == 0) {
            S.Dx()[koff + ia] = x;
            S.Dv()[koff + ia] = vy;
            S.Dc()[koff + ia] = c0;
            if (ia == 0) st->ev0w[iy] = ev;
        } else {
            P.line(g, wc, iy, 4)[1 + ia] = x;
            P.line(g, wc, iy, 5)[1 + ia] = c0;
            P.line(g, wc, iy, 6)[1 + ia] = vy;
            if (ia == 0) st->ev0r[iy] = ev;
        }
    }
}

// ---- envelope, tail and store of period t for one income state: one CTA per (state, model) -------------------
__global__ void __launch_bounds__(kEnvThreadsG) bigg_env_kernel(DeviceBatch B, unsigned char* scratch, size_t stride, int t) {
    const int iy = blockIdx.x, lm = blockIdx.y;
    const int m = B.first + lm;
    const int tid = threadIdx.x;
    __shared__ DevModel s_p;
    __shared__ int s_n[2];
    __shared__ int s_flag;
    if (tid == 0) {
        s_p = B.models[m];
        s_flag = 0;
    }
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny;
    const BigGPlan P(B.lay.cap, B.lay.isec_cap, ny);
    unsigned char* g = scratch + (size_t)lm * stride;
    BigGState* st = P.state(g);
    EnvScratchG S = P.scratch(g, iy);
    const int w = t & 1;
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const int koff = na / 10 - 1 > 0 ? na / 10 - 1 : 0;
    const double a0 = avec[0];
    const double ev0w = st->ev0w[iy], ev0r = st->ev0r[iy];
    const size_t ss = state_slot(B.lay, m, t, iy);
    const bool first_period = t == 0;  // see solve_f_kernel
    double* wcx = P.line(g, w, iy, 0); double* wcy = P.line(g, w, iy, 1);
    double* wvx = P.line(g, w, iy, 2); double* wvy = P.line(g, w, iy, 3);
    double* rx = P.line(g, w, iy, 4);  double* rcy = P.line(g, w, iy, 5); double* rvy = P.line(g, w, iy, 6);
    if (B.dirty) {
        double2* d = B.dirty + line_slot(B.lay, m, t, iy, 0, 0) * (size_t)na;
        for (int k = tid; k < na; k += kEnvThreadsG) {
            d[k] = make_double2(S.Dx()[koff + k], S.Dc()[koff + k]);
            d[(size_t)na + k] = make_double2(S.Dx()[koff + k], S.Dv()[koff + k]);
            d[(size_t)2 * na + k] = make_double2(rx[1 + k], rcy[1 + k]);
            d[(size_t)3 * na + k] = make_double2(rx[1 + k], rvy[1 + k]);
        }
    }
    worker_envelope<EnvScratchG, kEnvThreadsG>(S, koff, na, a0, ev0w, p, wvx, wvy, &s_n[0], wcx, wcy, &s_n[1], B.meta + ss,
                                               B.isects + ss * (size_t)B.lay.isec_cap, B.removed + ss * (size_t)B.lay.rem_words,
                                               B.lay.rem_words, first_period);
    const int n_wv = s_n[0], n_wc = s_n[1], n_r = na + 1;
    if (tid == 0) {
        rx[0] = a0;
        rcy[0] = 0.0;
        rvy[0] = ev0r;
    }
    __syncthreads();
    int flag = 0;
    for (int k = tid; k < n_r; k += kEnvThreadsG) {
        if (k + 1 < n_r) flag |= knot_pair_flag(rx[k], rx[k + 1], first_period);
        if (!isfinite(rx[k]) || !isfinite(rcy[k]) || !isfinite(rvy[k])) flag |= DCEGM_ST_NONFINITE;
    }
    if (flag) atomicOr(&s_flag, flag);
    if (tid == 0 && S.cnt()[2]) atomicOr(&s_flag, S.cnt()[2]);
    store_line_big<kEnvThreadsG>(B, m, t, iy, 0, DCEGM_KIND_C, wcx, wcy, n_wc);
    store_line_big<kEnvThreadsG>(B, m, t, iy, 0, DCEGM_KIND_V, wvx, wvy, n_wv);
    store_line_big<kEnvThreadsG>(B, m, t, iy, 1, DCEGM_KIND_C, rx, rcy, n_r);
    store_line_big<kEnvThreadsG>(B, m, t, iy, 1, DCEGM_KIND_V, rx, rvy, n_r);
    __syncthreads();
    if (tid == 0) {
        B.vbound[ss * 2 + 0] = ev0w;
        B.vbound[ss * 2 + 1] = ev0r;
        const dcegm_env_meta mt = B.meta[ss];
        // the ny CTAs of a model add their integer counts with atomics (order does not matter for integers)
        if (mt.n_removed > 0) atomicAdd(&st->sum_envp, 1);
        atomicAdd(&st->sum_seg, mt.n_segments);
        atomicAdd(&st->sum_is, mt.n_isects);
        atomicAdd(&st->sum_rm, mt.n_removed);
        int ml = n_wc > n_wv ? n_wc : n_wv;
        ml = ml > n_r ? ml : n_r;
        atomicMax(&st->sum_maxlen, ml);
        atomicAdd(&st->sum_pts, (unsigned long long)(n_wc + n_wv + 2 * n_r));
        if (s_flag) atomicOr(&st->status, s_flag);
        st->n_wc[w][iy] = n_wc;
        st->n_wv[w][iy] = n_wv;
        st->n_r[w][iy] = n_r;
        st->vb_w[w][iy] = ev0w;
        st->vb_r[w][iy] = ev0r;
    }
}

// ---- per-model summary (same arithmetic and summation order as the tail of solve_g_kernel) --------------------
__global__ void bigg_summary_kernel(DeviceBatch B, unsigned char* scratch, size_t stride) {
    const int m = B.first + blockIdx.x;
    if (threadIdx.x != 0) return;
    const DevModel p = B.models[m];
    const int ny = p.ny;
    const BigGPlan P(B.lay.cap, B.lay.isec_cap, ny);
    unsigned char* g = scratch + (size_t)blockIdx.x * stride;
    const BigGState* st = P.state(g);
    const int w = 0;  // period 1 (t = 0) lives in set 0
    dcegm_summary sm;
    memset(&sm, 0, sizeof sm);
    sm.status = st->status;
    sm.n_env_periods = st->sum_envp;
    sm.tot_segments = st->sum_seg;
    sm.tot_isects = st->sum_is;
    sm.tot_removed = st->sum_rm;
    sm.max_len = st->sum_maxlen;
    sm.tot_points = (long long)st->sum_pts;
    const double fr[DCEGM_SUMMARY_NPROBE] = {0.02, 0.05, 0.1, 0.2, 0.3, 0.5, 0.7, 0.9};
    for (int q = 0; q < DCEGM_SUMMARY_NPROBE; q++) {
        const double mq = p.a_low + fr[q] * (p.a_high - p.a_low);
        double c = interp_y<kTopG>(P.line(g, w, 0, 0), P.line(g, w, 0, 1), st->n_wc[w][0], mq);
        c = c < p.cfloor ? p.cfloor : c;
        sm.c1_probe[q] = c;
        sm.v1_probe[q] = (mq < P.line(g, w, 0, 2)[1]) ? u_work(c, true, p) + p.beta * st->vb_w[w][0]
                                                      : interp_y<kTopG>(P.line(g, w, 0, 2), P.line(g, w, 0, 3), st->n_wv[w][0], mq);
    }
    double cs_c = 0.0, cs_v = 0.0;
    for (int iy = 0; iy < ny; iy++) {
        const double* wcy = P.line(g, w, iy, 1);
        const double* rcy = P.line(g, w, iy, 5);
        const double* wvy = P.line(g, w, iy, 3);
        const double* rvy = P.line(g, w, iy, 6);
        for (int k = 0; k < st->n_wc[w][iy]; k++) cs_c += wcy[k];
        for (int k = 0; k < st->n_r[w][iy]; k++) cs_c += rcy[k];
        for (int k = 0; k < st->n_wv[w][iy]; k++) cs_v += isfinite(wvy[k]) ? wvy[k] : 0.0;
        for (int k = 0; k < st->n_r[w][iy]; k++) cs_v += isfinite(rvy[k]) ? rvy[k] : 0.0;
    }
    sm.checksum_c = cs_c;
    sm.checksum_v = cs_v;
    B.summaries[m] = sm;
    B.status[m] = st->status;
}

// the whole backward recursion for models [first, first + count); `scratch` holds `count` regions of
// solver_g_big_scratch_bytes().  Returns the number of launches in *nlaunch.
cudaError_t launch_solve_g_big(const DeviceBatch& B0, int first, int count, unsigned char* scratch, cudaStream_t stream,
                               int* nlaunch) {
    DeviceBatch B = B0;
    B.first = first;
    B.order = nullptr;
    const int nT = B.lay.nT, na = B.lay.na, ny = B.lay.ny;
    const size_t stride = BigGPlan(B.lay.cap, B.lay.isec_cap, ny).bytes;
    int nl = 0;
    bigg_terminal_kernel<<<count, kThreads, 0, stream>>>(B, scratch, stride);
    nl++;
    for (int t = nT - 2; t >= 0; t--) {
        // the grid's y and z extents are limited to 65535: models go in slabs
        for (int c0 = 0; c0 < count; c0 += 32768) {
            const int cnt = count - c0 < 32768 ? count - c0 : 32768;
            DeviceBatch Bs = B;
            Bs.first = first + c0;
            unsigned char* scr = scratch + (size_t)c0 * stride;
            bigg_egm_kernel<<<dim3((na + kThreads - 1) / kThreads, ny, cnt), kThreads, 0, stream>>>(Bs, scr, stride, t);
            bigg_env_kernel<<<dim3(ny, cnt), kEnvThreadsG, 0, stream>>>(Bs, scr, stride, t);
            nl += 2;
        }
    }
    bigg_summary_kernel<<<count, 32, 0, stream>>>(B, scratch, stride);
    nl++;
    if (nlaunch) *nlaunch = nl;
    return cudaGetLastError();
}

}  // namespace dcegm
