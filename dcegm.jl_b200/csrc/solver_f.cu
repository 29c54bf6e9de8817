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

#include "device_common.cuh"
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

// One set of the four polylines of a period: worker policy / value (own knots each), retiree policy and
// value (shared knots: both are (a + c0) with the same prepended point).  Offsets into dc_smem only.
struct LineSet {
    int base, cap8;
    int n_wc, n_wv, n_r;
    double vb_w, vb_r;  // Envelope.vbound of the two value lines
    __device__ __forceinline__ double* arr(int a) const { return reinterpret_cast<double*>(dc_smem + base + a * cap8); }
    __device__ __forceinline__ double* wc_x() const { return arr(0); }
    __device__ __forceinline__ double* wc_y() const { return arr(1); }
    __device__ __forceinline__ double* wv_x() const { return arr(2); }
    __device__ __forceinline__ double* wv_y() const { return arr(3); }
    __device__ __forceinline__ double* r_x() const { return arr(4); }
    __device__ __forceinline__ double* rc_y() const { return arr(5); }
    __device__ __forceinline__ double* rv_y() const { return arr(6); }
};

// one line -> HBM (interleaved x,y), each point written once
__device__ inline void store_line(const DeviceBatch& B, int m, int it, int iy, int id, int kind, const double* x,
                                  const double* y, int n) {
    const size_t slot = line_slot(B.lay, m, it, iy, id, kind);
    if (threadIdx.x == 0) B.lens[slot] = n;
    if (!B.lines) return;
    double2* dst = B.lines + slot * (size_t)B.lay.cap;
    const int nn = n < B.lay.cap ? n : B.lay.cap;
    for (int k = threadIdx.x; k < nn; k += kThreads) dst[k] = make_double2(x[k], y[k]);
}

// vfun (dc_algo.jl:220-251) for the value line (vx, vy, n) with bound vb
__device__ __forceinline__ double vfun_dev(bool worker_line, bool next_terminal, double c1, double m1, const double* vx,
                                           const double* vy, int n, double vb, const DevModel& p) {
    if (next_terminal && p.nu > 0) return p.nu * u_plain(c1 + p.bbar, p) - 0.0;  // bequest(c1,false,p)
    const bool mask = next_terminal || (m1 < vx[1]);
    if (mask) return u_work(c1, worker_line, p) + p.beta * vb;
    return interp_y(vx, vy, n, m1);
}

struct NodeOut {
    double mu, ls, v2;
};

// One quadrature node of the EGM step for ANY input values (dc_algo.jl:33-53): binary searches and
// the general pow.  Used when the fast path below does not apply (bequest period, delta > 0 for the
// retiree, cfloor <= 0, non-monotone or non-finite next-period lines, exponents that are not finite).
static __device__ __noinline__ void egm_node_generic(const LineSet* Np, const DevModel* pp, double m1, int id, bool next_terminal,
                                              bool beq, NodeOut* out) {
    const LineSet& N = *Np;
    const DevModel& p = *pp;
    const bool can_skip = p.cfloor > 0.0;  // 0*up(c) == 0 exactly when c >= cfloor > 0
    const bool need_w = (id == 0) || (p.delta != 0.0) || !can_skip;
    // future retiree (iid = 2): policy and value share their knots
    const int ir = left_knot(N.r_x(), N.n_r, m1);
    double c2 = interp_at(N.r_x(), N.rc_y(), N.n_r, ir, m1);
    c2 = c2 < p.cfloor ? p.cfloor : c2;  // floory! (line.jl:131-139)
    double v2;
    if (next_terminal && p.nu > 0)
        v2 = p.nu * u_plain(c2 + p.bbar, p) - 0.0;
    else if (next_terminal || m1 < N.r_x()[1])
        v2 = u_work(c2, false, p) + p.beta * N.vb_r;
    else
        v2 = interp_at(N.r_x(), N.rv_y(), N.n_r, ir, m1);
    double c1 = 0.0, v1 = 0.0;
    if (need_w) {
        c1 = interp_y(N.wc_x(), N.wc_y(), N.n_wc, m1);
        c1 = c1 < p.cfloor ? p.cfloor : c1;
        v1 = vfun_dev(true, next_terminal, c1, m1, N.wv_x(), N.wv_y(), N.n_wv, N.vb_w, p);
    }
    double pw, ls = 0.0;
    if (id == 0 || p.delta != 0.0) {
        // ccp / logsum (dc_algo.jl:316-332): the larger value's exp is exp(0) = 1
        const bool w_hi = v1 > v2;
        const double mx = w_hi ? v1 : v2;
        const double d = ((w_hi ? v2 : v1) - mx) / p.lambda;
        const double e = dc_exp(d);  // exactly 0 below -745.2
        const double e1 = w_hi ? 1.0 : e, e2 = w_hi ? e : 1.0;
        const double s = e1 + e2;
        pw = e1 / s;
        ls = mx + p.lambda * dc_log(s);
        if (id == 1) pw = p.delta;  // dc_algo.jl:44
    } else {
        pw = p.delta;  // == 0
    }
    if (beq) {  // dc_algo.jl:46-48
        c1 = c1 + p.bbar;
        c2 = c2 + p.bbar;
    }
    double mu;
    if (can_skip && pw == 1.0)
        mu = up(c1, p);
    else if (can_skip && pw == 0.0)
        mu = up(c2, p);
    else
        mu = pw * up(c1, p) + (1.0 - pw) * up(c2, p);
    out->mu = mu;
    out->ls = ls;
    out->v2 = v2;
}

static __device__ __noinline__ double dc_pow_cold(double x, double y) { return dc_pow(x, y); }
// x^y; straight-line code for positive normal x (y finite), the general routine otherwise
__device__ __forceinline__ double pow_hot(double x, double y) {
    if (!dc_is_norm_pos(x)) return dc_pow_cold(x, y);
    double lh, ll;
    dc_log_parts_pos(x, &lh, &ll);
    return dc_pow_from_log(lh, ll, y);
}
// two powers with one exponent, evaluated together (independent chains interleave); one cold branch
__device__ __forceinline__ void pow_hot2(double x1, double x2, double y, double* r1, double* r2) {
    const bool ok = dc_is_norm_pos(x1) && dc_is_norm_pos(x2);
    double h1, l1, h2, l2;
    dc_log_parts_pos(ok ? x1 : 1.0, &h1, &l1);
    dc_log_parts_pos(ok ? x2 : 1.0, &h2, &l2);
    *r1 = dc_pow_from_log(h1, l1, y);
    *r2 = dc_pow_from_log(h2, l2, y);
    if (!ok) {
        *r1 = dc_pow_cold(x1, y);
        *r2 = dc_pow_cold(x2, y);
    }
}
// utils.jl:40-46 / 72-78 from ln c = lh + ll (c positive normal)
__device__ __forceinline__ double u_from_log(double lh, double ll, bool working, const DevModel& p) {
    const double dis = working ? p.alpha : 0.0;
    if (p.gamma_is_one) return (lh + ll) - dis;
    return p.oneover_oneminusgamma * (dc_pow_from_log(lh, ll, p.oneminusgamma) - 1.0) - dis;
}
__device__ __forceinline__ double up_from_log(double lh, double ll, double c, const DevModel& p) {
    if (p.gamma_is_one) return 1.0 / c;
    return dc_pow_from_log(lh, ll, p.neg_gamma);
}
// interp (line.jl:220-254), y only, on a strictly increasing finite line, branch-free.  cnt = number
// of knots <= q.  Interior: f = (q - x0)/(x1 - x0), (1-f) y0 + f y1.  Outside the knots Julia's
// extrapolate(.., Line()) is y_end + ((y1 - y0)/(x1 - x0)) (q - x_end) on the first / last interval: the
// same divisor, another numerator — so one reciprocal serves both forms and the choice is a select.
struct KnotCell {
    int i;             // left knot of the interval used
    double x0, x1, r;  // its abscissae, r = 1/(x1 - x0)
    double dq;         // q - x0 (interior, left of the line) or q - x1 (right of the line)
    bool lo, ext;      // q < x[0]; q outside [x[0], x[n-1]]
};
__device__ __forceinline__ KnotCell knot_cell(const double* __restrict__ x, int n, int cnt, double q) {
    KnotCell c;
    int i = cnt - 1;
    i = i < 0 ? 0 : i;
    i = i > n - 2 ? n - 2 : i;
    c.i = i;
    c.x0 = x[i];
    c.x1 = x[i + 1];
    c.lo = cnt == 0;
    const bool hi = q > c.x1;  // only possible on the last interval
    c.ext = c.lo | hi;
    c.dq = q - (hi ? c.x1 : c.x0);
    c.r = dc_rcp_fast(c.x1 - c.x0);
    return c;
}
__device__ __forceinline__ double interp_cell(const KnotCell& c, const double* __restrict__ y) {
    const double y0 = y[c.i], y1 = y[c.i + 1];
    const double d = c.x1 - c.x0;
    const double t = dc_div_with_rcp(c.ext ? y1 - y0 : c.dq, d, c.r);
    const double inner = (1.0 - t) * y0 + t * y1;
    const double outer = (c.lo ? y0 : y1) + t * c.dq;
    return c.ext ? outer : inner;
}
__device__ __forceinline__ double interp_cnt(const double* __restrict__ x, const double* __restrict__ y, int n, int cnt,
                                             double q) {
    return interp_cell(knot_cell(x, n, cnt, q), y);
}

#ifdef DCEGM_PHASE_CLOCKS
__device__ unsigned long long dc_phase_total[kPhaseSlots];
#endif

__global__ void __launch_bounds__(kThreads, kCtasPerSm) solve_f_kernel(DeviceBatch B) {
    const int m = B.first + blockIdx.x;
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
    const int nbuck = knot_index_buckets(2 * P.cap8, 3);
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
        bool fast = model_fast && !beq && !(s_status & (DCEGM_ST_NONMONOTONE | DCEGM_ST_NONFINITE)) && N.n_wc >= 2 &&
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
            const int nmax = max(N.n_r, max(N.n_wc, N.n_wv));
            for (int k = tid; k < nmax; k += kThreads) {  // the three tables in one pass
                if (k < N.n_r) knot_index_build_step(Kr, N.r_x(), N.n_r, nbuck, k);
                if (k < N.n_wc) knot_index_build_step(Kc, N.wc_x(), N.n_wc, nbuck, k);
                if (k < N.n_wv) knot_index_build_step(Kv, N.wv_x(), N.n_wv, nbuck, k);
            }
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
                const double aR = a * p.R;
                double acc_mu = 0.0, acc_ls = 0.0;
#pragma unroll 1
                for (int iy = 0; iy < ny; iy++) {
                    const double m1 = aR + s_off[iy];
                    double mu, ls;
                    bool redo = !fast;
                    if (fast) {
                        // future retiree (iid = 2): policy and value share their knots
                        const int nr = count_le_indexed(Kr, rX, n_r, m1);
                        const int nc = count_le_indexed(Kc, cX, n_wc, m1);
                        const KnotCell kr = knot_cell(rX, n_r, nr, m1);
                        double c2 = interp_cell(kr, rC);
                        double c1 = interp_cnt(cX, cY, n_wc, nc, m1);
                        c2 = c2 < p.cfloor ? p.cfloor : c2;  // floory! (line.jl:131-139)
                        c1 = c1 < p.cfloor ? p.cfloor : c1;
                        redo = !(dc_is_norm_pos(c1) && dc_is_norm_pos(c2));
                        double l1h, l1l, l2h, l2l;
                        dc_log_parts_pos(c1, &l1h, &l1l);
                        dc_log_parts_pos(c2, &l2h, &l2l);
                        // vfun (dc_algo.jl:220-251): analytic in the credit-constrained region, else interpolated
                        const bool mask2 = next_terminal || m1 < rx1, mask1 = next_terminal || m1 < wvx1;
                        double v1 = 0.0, v2 = 0.0;
                        if (mask1 || mask2) {  // both evaluated together: two independent chains
                            v1 = u_from_log(l1h, l1l, true, p) + bvb_w;
                            v2 = u_from_log(l2h, l2l, false, p) + bvb_r;
                        }
                        if (!mask2) v2 = interp_cell(kr, rV);
                        if (!mask1) v1 = interp_cnt(vX, vY, n_wv, count_le_indexed(Kv, vX, n_wv, m1), m1);
                        // ccp / logsum (dc_algo.jl:316-332): the larger value's exp is exp(0) = 1
                        const bool w_hi = v1 > v2;
                        const double mx = w_hi ? v1 : v2;
                        const double d = dc_div_fast((w_hi ? v2 : v1) - mx, p.lambda);
                        // marginal utility of the preferred choice
                        const double u_hi = up_from_log(w_hi ? l1h : l2h, w_hi ? l1l : l2l, w_hi ? c1 : c2, p);
                        if (can_skip && d < -745.2) {
                            // the loser's exp underflows to exactly 0: P(work) is 0 or 1, log(1 + 0) = 0
                            ls = mx + p.lambda * 0.0;
                            mu = u_hi;
                        } else {
                            const double e = dc_exp(d);
                            const double u_lo = up_from_log(w_hi ? l2h : l1h, w_hi ? l2l : l1l, w_hi ? c2 : c1, p);
                            const double e1 = w_hi ? 1.0 : e, e2 = w_hi ? e : 1.0;
                            const double s = e1 + e2;
                            const double pw = dc_div_fast(e1, s);  // s in [1, 2]
                            redo |= !dc_is_norm_pos(s);
                            double sh, sl;
                            dc_log_parts_pos(s, &sh, &sl);
                            ls = mx + p.lambda * (sh + sl);
                            const double u1 = w_hi ? u_hi : u_lo, u2 = w_hi ? u_lo : u_hi;
                            mu = pw * u1 + (1.0 - pw) * u2;
                            mu = (can_skip && pw == 1.0) ? u1 : mu;
                            mu = (can_skip && pw == 0.0) ? u2 : mu;
                        }
                    }
                    if (redo) {
                        const LineSet Ncopy = N;
                        NodeOut o;
                        egm_node_generic(&Ncopy, &s_p, m1, 0, next_terminal, beq, &o);
                        mu = o.mu;
                        ls = o.ls;
                    }
                    const double w = s_wgt[iy];
                    // quadrature sums, ascending node order (dc_algo.jl:58,74)
                    acc_mu = iy == 0 ? w * mu : acc_mu + w * mu;
                    acc_ls = iy == 0 ? w * ls : acc_ls + w * ls;
                }
                // ---- retiree today (id = 1) at the same grid point ------------------------------------
                double acc_mu_r = 0.0, acc_ls_r = 0.0, acc_v2 = 0.0;
                {
                    double mu = 0.0, ls = 0.0, v2 = 0.0;
#pragma unroll 1
                    for (int iy = 0; iy < ny; iy++) {
                        if (iy == 0 || !same) {
                            const double m1 = aR + s_off[ny + iy];
                            bool redo = !fast_r;
                            if (fast_r) {
                                const int nr = count_le_indexed(Kr, rX, n_r, m1);
                                const KnotCell kr = knot_cell(rX, n_r, nr, m1);
                                double c2 = interp_cell(kr, rC);
                                c2 = c2 < p.cfloor ? p.cfloor : c2;
                                redo = !dc_is_norm_pos(c2);
                                double l2h, l2l;
                                dc_log_parts_pos(c2, &l2h, &l2l);
                                if (next_terminal || m1 < rx1)
                                    v2 = u_from_log(l2h, l2l, false, p) + bvb_r;
                                else
                                    v2 = interp_cell(kr, rV);
                                mu = up_from_log(l2h, l2l, c2, p);  // P(work) = delta = 0 (dc_algo.jl:44)
                            }
                            if (redo) {
                                const LineSet Ncopy = N;
                                NodeOut o;
                                egm_node_generic(&Ncopy, &s_p, m1, 1, next_terminal, beq, &o);
                                mu = o.mu;
                                ls = o.ls;
                                v2 = o.v2;
                            }
                        }
                        const double w = s_wgt[iy];
                        acc_mu_r = iy == 0 ? w * mu : acc_mu_r + w * mu;
                        acc_ls_r = iy == 0 ? w * ls : acc_ls_r + w * ls;
                        acc_v2 = iy == 0 ? w * v2 : acc_v2 + w * v2;
                    }
                }
                // ---- Euler inversion and today's values, both choices together (dc_algo.jl:50-78) -----
                const double rhs_w = scale * acc_mu, rhs_r = scale * acc_mu_r;
                double c0_w, c0_r, pw_w, pw_r;  // iup (dc_algo.jl:62), c0^(1-gamma)
                if (p.gamma_is_one) {
                    c0_w = 1.0 / rhs_w;
                    c0_r = 1.0 / rhs_r;
                    pw_w = pw_r = 0.0;
                } else {
                    pow_hot2(rhs_w, rhs_r, p.neg_oneover_gamma, &c0_w, &c0_r);
                    pow_hot2(c0_w, c0_r, p.oneminusgamma, &pw_w, &pw_r);
                }
                double uc_w, uc_r;  // u(c0, working), utils.jl:40-46
                if (p.gamma_is_one) {
                    uc_w = dc_log(c0_w) - p.alpha;
                    uc_r = dc_log(c0_r) - 0.0;
                } else {
                    uc_w = p.oneover_oneminusgamma * (pw_w - 1.0) - p.alpha;
                    uc_r = p.oneover_oneminusgamma * (pw_r - 1.0) - 0.0;
                }
                double ev;
                if (p.delta != 0.0)
                    ev = (1.0 - p.delta) * acc_v2 + p.delta * acc_ls_r;  // dc_algo.jl:76
                else
                    ev = acc_v2;
                S.Dx()[koff + ia] = a + c0_w;
                S.Dv()[koff + ia] = uc_w + p.beta * acc_ls;  // dc_algo.jl:74,78
                S.Dc()[koff + ia] = c0_w;
                C.r_x()[1 + ia] = a + c0_r;
                C.rc_y()[1 + ia] = c0_r;
                C.rv_y()[1 + ia] = uc_r + p.beta * ev;  // dc_algo.jl:78
                if (ia == 0) {
                    s_ev0[0] = acc_ls;
                    s_ev0[1] = ev;
                }
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
        worker_envelope(S, koff, na, a0, ev0w, p, C.wv_x(), C.wv_y(), &s_n[0], C.wc_x(), C.wc_y(), &s_n[1], B.meta + ss,
                        B.isects + ss * (size_t)B.lay.isec_cap, B.removed + ss * (size_t)B.lay.rem_words,
                        B.lay.rem_words);
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
            if (k + 1 < na + 1 && !(C.r_x()[k + 1] > C.r_x()[k])) flag |= DCEGM_ST_NONMONOTONE;
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
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(solve_f_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    solve_f_kernel<<<count, kThreads, smem, stream>>>(B);
    return cudaGetLastError();
}

}  // namespace dcegm
