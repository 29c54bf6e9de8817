// egm_node.cuh — one quadrature node of the retirement-model EGM step (dc_algo.jl:33-53) and the polyline
// set of one period, shared by the persistent shared-memory solver (solver_f.cu) and the fine-grid solver
// (solver_f_big.cu).  Both evaluate exactly the same sequence of IEEE operations per node, so their
// results are bit-identical to each other and to the oracle's portable mode.
#pragma once
#include <math_constants.h>

#include "device_common.cuh"

namespace dcegm {

// One set of the four polylines of a period: worker policy / value (own knots each), retiree policy and
// value (shared knots: both are (a + c0) with the same prepended point).  Offsets into the scratch memory only.
template <class Mem>
struct LineSetT : Mem {
    static constexpr int kTop = Mem::kTop;
    int base, cap8;
    int n_wc, n_wv, n_r;
    double vb_w, vb_r;  // Envelope.vbound of the two value lines
    __device__ __forceinline__ double* arr(int a) const { return reinterpret_cast<double*>(this->mem() + base + a * cap8); }
    __device__ __forceinline__ double* wc_x() const { return arr(0); }
    __device__ __forceinline__ double* wc_y() const { return arr(1); }
    __device__ __forceinline__ double* wv_x() const { return arr(2); }
    __device__ __forceinline__ double* wv_y() const { return arr(3); }
    __device__ __forceinline__ double* r_x() const { return arr(4); }
    __device__ __forceinline__ double* rc_y() const { return arr(5); }
    __device__ __forceinline__ double* rv_y() const { return arr(6); }
};
using LineSet = LineSetT<MemShared>;
using LineSetG = LineSetT<MemGlobal>;

// one line -> HBM (interleaved x,y), each point written once
template <int NT = kThreads>
__device__ inline void store_line(const DeviceBatch& B, int m, int it, int iy, int id, int kind, const double* x,
                                  const double* y, int n) {
    const size_t slot = line_slot(B.lay, m, it, iy, id, kind);
    if (threadIdx.x == 0) B.lens[slot] = n;
    if (!B.lines) return;
    double2* dst = B.lines + slot * (size_t)B.lay.cap;
    const int nn = n < B.lay.cap ? n : B.lay.cap;
    for (int k = threadIdx.x; k < nn; k += NT) dst[k] = make_double2(x[k], y[k]);
}

// vfun (dc_algo.jl:220-251) for the value line (vx, vy, n) with bound vb
template <int Top = kSearchTop>
__device__ __forceinline__ double vfun_dev(bool worker_line, bool next_terminal, double c1, double m1, const double* vx,
                                           const double* vy, int n, double vb, const DevModel& p) {
    if (next_terminal && p.nu > 0) return p.nu * u_plain(c1 + p.bbar, p) - 0.0;  // bequest(c1,false,p)
    const bool mask = next_terminal || (m1 < vx[1]);
    if (mask) return u_work(c1, worker_line, p) + p.beta * vb;
    return interp_y<Top>(vx, vy, n, m1);
}

struct NodeOut {
    double mu, ls, v2;
};

// One quadrature node of the EGM step for ANY input values (dc_algo.jl:33-53): binary searches and
// the general pow.  Used when the fast path below does not apply (bequest period, delta > 0 for the
// retiree, cfloor <= 0, non-monotone or non-finite next-period lines, exponents that are not finite).
template <class LS>
static __device__ __noinline__ void egm_node_generic(const LS* Np, const DevModel* pp, double m1, int id, bool next_terminal,
                                              bool beq, NodeOut* out) {
    const LS& N = *Np;
    constexpr int Top = LS::kTop;
    const DevModel& p = *pp;
    const bool can_skip = p.cfloor > 0.0;  // 0*up(c) == 0 exactly when c >= cfloor > 0
    const bool need_w = (id == 0) || (p.delta != 0.0) || !can_skip;
    // future retiree (iid = 2): policy and value share their knots
    const int ir = left_knot<Top>(N.r_x(), N.n_r, m1);
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
        c1 = interp_y<Top>(N.wc_x(), N.wc_y(), N.n_wc, m1);
        c1 = c1 < p.cfloor ? p.cfloor : c1;
        v1 = vfun_dev<Top>(true, next_terminal, c1, m1, N.wv_x(), N.wv_y(), N.n_wv, N.vb_w, p);
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
    dc_log_parts_pos2(ok ? x1 : 1.0, ok ? x2 : 1.0, &h1, &l1, &h2, &l2);
    dc_pow_from_log2(h1, l1, h2, l2, y, r1, r2);
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
// u(c1, working) and u(c2, not working) together
__device__ __forceinline__ void u_from_log2(double l1h, double l1l, double l2h, double l2l, const DevModel& p, double* u1,
                                            double* u2) {
    if (p.gamma_is_one) {
        *u1 = (l1h + l1l) - p.alpha;
        *u2 = (l2h + l2l) - 0.0;
        return;
    }
    double p1, p2;
    dc_pow_from_log2(l1h, l1l, l2h, l2l, p.oneminusgamma, &p1, &p2);
    *u1 = p.oneover_oneminusgamma * (p1 - 1.0) - p.alpha;
    *u2 = p.oneover_oneminusgamma * (p2 - 1.0) - 0.0;
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

}  // namespace dcegm
