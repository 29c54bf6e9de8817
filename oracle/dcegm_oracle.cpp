// dcegm_oracle.cpp — CPU restatement of the DCEGM.jl hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under dcegm.jl_b200/ may include, link or call this
// file; it is the checker for the CUDA path (tests/, __graft_entry__.smoke(), and the
// cpu_baseline / --impl reference legs of bench.py).
//
// Parity status: PINNED against the reference's own golden data for the FModel path
// (matlab/output/*.csv through test/F_test.jl's checks, see tests/test_oracle_golden.py)
// and against the known answers of test/Envtests.jl and test/basics.jl.  GModel,
// minimal_EGM, bequest (nu>0) and delta>0 branches have no reference test: for those the
// oracle is "parity unpinned"  and only restates the Julia source.
//
// The reference's Julia implementation cannot run here (no julia binary); its third-party
// arithmetic is restated from the published definitions:
//   Interpolations.jl  Gridded(Linear()) + extrapolate(.., Line())   -> interp_pt()
//   Roots.jl           fzero(f,a,b) (bracketing)                      -> closed-form root of the
//                      difference of two linear pieces; a bisection variant (literal=1) is kept
//                      for cross-checking
//   Distributions.jl   quantile(Normal(0,1), p)                       -> norminv()
//   FastGaussQuadrature.jl gausshermite(n)                            -> gauss_hermite()
// All four are unpinned in the reference (no Manifest.toml).
//
// Every function cites the reference file:line (relative to the reference tree) it follows.
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -pthread).
#include "../include/dcegm.h"
#include "../dcegm.jl_b200/csrc/dc_math.h"  // shared reproducible math (portable mode only)
#include "../dcegm.jl_b200/csrc/dc_rng.h"   // shared definition of the simulator's counter-based shocks

#include <algorithm>
#include <atomic>
#include <thread>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

namespace {

struct Pt {
    double x, y;
};
typedef std::vector<Pt> Line;

struct Par : dcegm_param {
    double neg_gamma, oneminusgamma, oneover_oneminusgamma, neg_oneover_gamma;
    // 0: glibc libm (default; the mode pinned against the golden files)
    // 1: the reproducible pow/exp/log of dcegm.jl_b200/csrc/dc_math.h, which the CUDA kernels use too.
    //    In this mode GPU and oracle must agree bit for bit; the difference between the two modes
    //    is the reference algorithm's own sensitivity to 1-ulp libm differences.
    int portable;
};
inline double m_pow(double x, double y, int portable) { return portable ? dc_pow(x, y) : std::pow(x, y); }
inline double m_exp(double x, int portable) { return portable ? dc_exp(x) : std::exp(x); }
inline double m_log(double x, int portable) { return portable ? dc_log(x) : std::log(x); }

// param.jl:90-93
Par derive(const dcegm_param& p, int portable = 0) {
    Par q;
    q.portable = portable;
    static_cast<dcegm_param&>(q) = p;
    q.neg_gamma = (-1.0) * p.gamma;
    q.oneminusgamma = 1.0 - p.gamma;
    q.oneover_oneminusgamma = 1.0 / q.oneminusgamma;
    q.neg_oneover_gamma = (-1.0) / p.gamma;
    return q;
}

const double NEG_INF = -std::numeric_limits<double>::infinity();

// ---------------------------------------------------------------------------------------
// scalar economics — utils.jl
// ---------------------------------------------------------------------------------------
// utils.jl:23-29
double u_plain(double x, const Par& p) {
    if (p.gamma == 1.0) return m_log(x, p.portable);
    return p.oneover_oneminusgamma * m_pow(x, p.oneminusgamma, p.portable);
}
// utils.jl:40-46
double u_work(double x, bool working, const Par& p) {
    double dis = working ? p.alpha : 0.0;
    if (p.gamma == 1.0) return m_log(x, p.portable) - dis;
    return p.oneover_oneminusgamma * (m_pow(x, p.oneminusgamma, p.portable) - 1.0) - dis;
}
// utils.jl:6-9, 57-59
double bequest(double b, const Par& p) { return p.nu * u_plain(b + p.bbar, p); }
double bequest_work(double b, bool working, const Par& p) {
    return bequest(b, p) - (working ? p.alpha : 0.0);
}
// utils.jl:72-78
double up(double c, const Par& p) {
    if (p.gamma == 1.0) return 1.0 / c;
    return m_pow(c, p.neg_gamma, p.portable);
}
// utils.jl:102-108
double iup(double u, const Par& p) {
    if (p.gamma == 1.0) return 1.0 / u;
    return m_pow(u, p.neg_oneover_gamma, p.portable);
}
// utils.jl:243-250
double income(int it /*1-based*/, const Par& p, double shock) {
    int age = it + 19;
    if (it > p.retage) return p.pension;
    return std::exp(p.inc0 + p.inc1 * age - p.inc2 * (double)(age * age) + shock);
}

// ---------------------------------------------------------------------------------------
// host-side numerics — grids and quadrature
// ---------------------------------------------------------------------------------------
// Julia `range(lo, stop=hi, length=n)` builds a StepRangeLen in twice-precision arithmetic,
// i.e. each element is the (nearly) correctly rounded value of lo + i*(hi-lo)/(n-1).
// Restated with x87 long double (64-bit significand).   param.jl:205, dc_algo.jl:406
void linrange(double lo, double hi, int n, std::vector<double>& out) {
    out.resize(n);
    if (n == 1) {
        out[0] = lo;
        return;
    }
    long double l = lo, h = hi;
    for (int i = 0; i < n; i++) {
        long double v = (l * (long double)(n - 1 - i) + h * (long double)i) / (long double)(n - 1);
        out[i] = (double)v;
    }
    out[0] = lo;
    out[n - 1] = hi;
}

// utils.jl:261-297 — Gauss–Legendre nodes/weights on [lbnd, ubnd], literal port.
void quadpoints(int n, double lbnd, double ubnd, std::vector<double>& x, std::vector<double>& w) {
    x.assign(n, 0.0);
    w.assign(n, 0.0);
    const double EPS = 3.e-14;
    double x2 = ubnd, x1 = lbnd;
    int m = (int)std::floor((n + 1) / 2.0);
    double xm = (x2 + x1) / 2, xl = (x2 - x1) / 2;
    double z1 = 1.e99, pp = 0, p1 = 0, p2 = 0, p3 = 0, z = 0;
    const double pi = 3.141592653589793;
    for (int i = 1; i <= m; i++) {
        z = std::cos(pi * (i - 0.25) / (n + 0.5));
        while (std::fabs(z - z1) > EPS) {
            p1 = 1;
            p2 = 0;
            for (int j = 1; j <= n; j++) {
                p3 = p2;
                p2 = p1;
                p1 = ((2 * j - 1) * z * p2 - (j - 1) * p3) / j;
            }
            pp = n * (z * p1 - p2) / (z * z - 1);
            z1 = z;
            z = z1 - p1 / pp;
        }
        x[i - 1] = xm - xl * z;
        x[n - i] = xm + xl * z;
        w[i - 1] = 2 * xl / ((1 - z * z) * pp * pp);
        w[n - i] = w[i - 1];
    }
}

// quantile(Normal(0,1), p) — param.jl:195-196.  Wichura AS241 (PPND16) followed by one
// Halley step on erfc, which brings the result to ~1 ulp.
double norminv(double p) {
    if (!(p > 0.0)) return NEG_INF;
    if (!(p < 1.0)) return -NEG_INF;
    double q = p - 0.5, r, val;
    if (std::fabs(q) <= 0.425) {
        r = 0.180625 - q * q;
        val = q *
              (((((((2.5090809287301226727e3 * r + 3.3430575583588128105e4) * r + 6.7265770927008700853e4) * r +
                   4.5921953931549871457e4) * r + 1.3731693765509461125e4) * r + 1.9715909503065514427e3) * r +
                1.3314166789178437745e2) * r + 3.3871328727963666080e0) /
              (((((((5.2264952788528545610e3 * r + 2.8729085735721942674e4) * r + 3.9307895800092710610e4) * r +
                   2.1213794301586595867e4) * r + 5.3941960214247511077e3) * r + 6.8718700749205790830e2) * r +
                4.2313330701600911252e1) * r + 1.0);
    } else {
        r = q < 0 ? p : 1.0 - p;
        r = std::sqrt(-std::log(r));
        if (r <= 5.0) {
            r -= 1.6;
            val = (((((((7.74545014278341407640e-4 * r + 2.27238449892691845833e-2) * r + 2.41780725177450611770e-1) * r +
                      1.27045825245236838258e0) * r + 3.64784832476320460504e0) * r + 5.76949722146069140550e0) * r +
                    4.63033784615654529590e0) * r + 1.42343711074968357734e0) /
                  (((((((1.05075007164441684324e-9 * r + 5.47593808499534494600e-4) * r + 1.51986665636164571966e-2) * r +
                      1.48103976427480074590e-1) * r + 6.89767334985100004550e-1) * r + 1.67638483018380384940e0) * r +
                    2.05319162663775882187e0) * r + 1.0);
        } else {
            r -= 5.0;
            val = (((((((2.01033439929228813265e-7 * r + 2.71155556874348757815e-5) * r + 1.24266094738807843860e-3) * r +
                      2.65321895265761230930e-2) * r + 2.96560571828504891230e-1) * r + 1.78482653991729133580e0) * r +
                    5.46378491116411436990e0) * r + 6.65790464350110377720e0) /
                  (((((((2.04426310338993978564e-15 * r + 1.42151175831644588870e-7) * r + 1.84631831751005468180e-5) * r +
                      7.86869131145613259100e-4) * r + 1.48753612908506148525e-2) * r + 1.36929880922735805310e-1) * r +
                    5.99832206555887937690e-1) * r + 1.0);
        }
        if (q < 0) val = -val;
    }
    // Halley refinement: e = Phi(val) - p
    double e = 0.5 * std::erfc(-val / std::sqrt(2.0)) - p;
    double uu = e * std::sqrt(2.0 * 3.141592653589793) * std::exp(val * val / 2.0);
    val = val - uu / (1.0 + val * uu / 2.0);
    return val;
}

// gausshermite(n) of FastGaussQuadrature.jl (physicists' weight exp(-x^2)) — param.jl:279,
// minimal_EGM.jl:4.  Newton iteration on the orthonormal recurrence in long double;
// nodes ascending.
void gauss_hermite(int n, std::vector<double>& x, std::vector<double>& w) {
    x.assign(n, 0.0);
    w.assign(n, 0.0);
    const long double PIM4 = 0.7511255444649425L;  // pi^(-1/4)
    int m = (n + 1) / 2;
    long double z = 0, pp = 0;
    std::vector<long double> roots(m);
    for (int i = 0; i < m; i++) {
        if (i == 0)
            z = sqrtl((long double)(2 * n + 1)) - 1.85575L * powl((long double)(2 * n + 1), -0.16667L);
        else if (i == 1)
            z -= 1.14L * powl((long double)n, 0.426L) / z;
        else if (i == 2)
            z = 1.86L * z - 0.86L * roots[0];
        else if (i == 3)
            z = 1.91L * z - 0.91L * roots[1];
        else
            z = 2.0L * z - roots[i - 2];
        for (int its = 0; its < 200; its++) {
            long double p1 = PIM4, p2 = 0.0L;
            for (int j = 0; j < n; j++) {
                long double p3 = p2;
                p2 = p1;
                p1 = z * sqrtl(2.0L / (j + 1)) * p2 - sqrtl((long double)j / (j + 1)) * p3;
            }
            pp = sqrtl((long double)(2 * n)) * p2;
            long double z1 = z;
            z = z1 - p1 / pp;
            if (fabsl(z - z1) <= 1e-19L * (1.0L + fabsl(z))) break;
        }
        roots[i] = z;
        // largest root first: ascending order puts it at the end
        x[n - 1 - i] = (double)z;
        x[i] = (double)(-z);
        w[i] = (double)(2.0L / (pp * pp));
        w[n - 1 - i] = w[i];
    }
    if (n % 2 == 1) x[n / 2] = 0.0;
}

// utils.jl:178-197
void rouwenhorst(double rho, double mu_eps, double sigma_eps, int n, std::vector<double>& z,
                 std::vector<double>& P /* row-major n x n */) {
    double q = (rho + 1) / 2;
    double nu = std::sqrt((n - 1) / (1 - rho * rho)) * sigma_eps;
    std::vector<double> cur = {q, 1 - q, 1 - q, q};
    int sz = 2;
    for (int i = 2; i <= n - 1; i++) {
        int ns = i + 1;
        std::vector<double> nxt(ns * ns, 0.0);
        auto at = [&](int r, int c) -> double {
            if (r < 0 || c < 0 || r >= sz || c >= sz) return 0.0;
            return cur[r * sz + c];
        };
        for (int r = 0; r < ns; r++)
            for (int c = 0; c < ns; c++) {
                double v = q * at(r, c) + (1 - q) * at(r, c - 1);
                v = v + (1 - q) * at(r - 1, c);
                v = v + q * at(r - 1, c - 1);
                nxt[r * ns + c] = v;
            }
        for (int r = 1; r <= i - 1; r++)
            for (int c = 0; c < ns; c++) nxt[r * ns + c] = nxt[r * ns + c] / 2;
        cur.swap(nxt);
        sz = ns;
    }
    P = cur;
    linrange(mu_eps / (1 - rho) - nu, mu_eps / (1 - rho) + nu, n, z);
}

// utils.jl:200-238
void scale_grid(double lb, double ub, int n, int logorder, std::vector<double>& out) {
    if (lb >= ub) lb = ub - std::fabs(ub) * 0.05;
    double off = 1;
    if (lb < 0) off = 1 - lb;
    if (logorder == 0) {
        linrange(lb, ub, n, out);
    } else if (logorder == 1) {
        linrange(std::log(lb + off), std::log(ub + off), n, out);
        for (auto& v : out) v = std::exp(v) - off;
    } else if (logorder == 2) {
        linrange(std::log(std::log(lb + off) + off), std::log(std::log(ub + off) + off), n, out);
        for (auto& v : out) v = std::exp(std::exp(v) - off) - off;
    } else {
        linrange(std::log(std::log(std::log(lb + off) + off) + off), std::log(std::log(std::log(ub + off) + off) + off), n, out);
        for (auto& v : out) v = std::exp(std::exp(std::exp(v) - off) - off) - off;
    }
}

// ---------------------------------------------------------------------------------------
// polyline primitives — line.jl
// ---------------------------------------------------------------------------------------
// searchsortedlast on the x coordinates, clamped to a valid left knot (0-based, [0,n-2])
inline int left_knot(const Pt* L, int n, double q) {
    int lo = 0, hi = n;  // number of knots with x <= q
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (L[mid].x <= q)
            lo = mid + 1;
        else
            hi = mid;
    }
    int i = lo - 1;
    if (i < 0) i = 0;
    if (i > n - 2) i = n - 2;
    return i;
}

// line.jl:220-254 with extrap=true: Gridded(Linear()) inside the knot range,
// extrapolate(...,Line()) outside.  Returns the whole Point (x is interpolated too,
// line.jl:236-250).
inline Pt interp_pt(const Pt* L, int n, double q) {
    Pt r;
    if (q < L[0].x) {
        double d = L[1].x - L[0].x;
        r.x = L[0].x + ((L[1].x - L[0].x) / d) * (q - L[0].x);
        r.y = L[0].y + ((L[1].y - L[0].y) / d) * (q - L[0].x);
        return r;
    }
    if (q > L[n - 1].x) {
        double d = L[n - 1].x - L[n - 2].x;
        r.x = L[n - 1].x + ((L[n - 1].x - L[n - 2].x) / d) * (q - L[n - 1].x);
        r.y = L[n - 1].y + ((L[n - 1].y - L[n - 2].y) / d) * (q - L[n - 1].x);
        return r;
    }
    int i = left_knot(L, n, q);
    double f = (q - L[i].x) / (L[i + 1].x - L[i].x);
    r.x = (1.0 - f) * L[i].x + f * L[i + 1].x;
    r.y = (1.0 - f) * L[i].y + f * L[i + 1].y;
    return r;
}
inline double interp_y(const Line& L, double q) { return interp_pt(L.data(), (int)L.size(), q).y; }

// line.jl:239-246 — extrap=false: -Inf outside the domain
inline Pt interp_pt_noextrap(const Line& L, double q) {
    if (q < L.front().x || q > L.back().x) return Pt{q, NEG_INF};
    return interp_pt(L.data(), (int)L.size(), q);
}

inline bool strictly_increasing(const Line& L) {
    for (size_t k = 1; k < L.size(); k++)
        if (!(L[k].x > L[k - 1].x)) return false;
    return true;
}
// status bits of a stored line whose knot vector interpolate() (line.jl:236) rejects; it = 1-based period of the line.
// Period-1 lines are never interpolated by the solve (dc_algo.jl:22-37 reads it+1 only): informational bit.
inline unsigned knot_flags(const Line& L, int it) {
    unsigned f = 0;
    for (size_t k = 1; k < L.size(); k++) {
        if (L[k].x > L[k - 1].x) continue;
        if (it == 1)
            f |= DCEGM_ST_NONMONO_FIRST;
        else
            f |= L[k].x == L[k - 1].x ? DCEGM_ST_DUP_KNOT : DCEGM_ST_NONMONOTONE;
    }
    return f;
}

// line.jl:40,348-351 — stable sort on x only
inline void sortx(Line& L) {
    std::stable_sort(L.begin(), L.end(), [](const Pt& a, const Pt& b) { return a.x < b.x; });
}

// ---------------------------------------------------------------------------------------
// envelope algebra — envelope.jl
// ---------------------------------------------------------------------------------------
struct Envelope {
    Line env;
    std::vector<Pt> isects;
    std::vector<int> removed;  // 0-based indices into the line given to secondary_envelope
    double vbound = 0.0;       // envelope.jl:37
    int n_segments = 1;
    int n_prepended = 0;
};

// envelope.jl:80-126 (+ splitat, line.jl:323-345).  Returns false if the reference would
// produce no section at all (an entirely non-rising line).
bool split_line(const Line& o, std::vector<Line>& out) {
    out.clear();
    int n = (int)o.size();
    std::vector<char> ii(n > 0 ? n - 1 : 0);
    bool all_rising = true;
    for (int k = 0; k + 1 < n; k++) {
        ii[k] = o[k + 1].x > o[k].x;
        if (!ii[k]) all_rising = false;
    }
    if (all_rising) {
        out.push_back(o);
        return true;
    }
    int start = 0;  // first point of the remaining line
    int cuts = 0;
    while (true) {
        // j = first direction index (relative to start) that differs from ii[start]
        int j = -1;
        for (int k = start + 1; k + 1 < n; k++)
            if (ii[k] != ii[start]) {
                j = k;
                break;
            }
        if (j < 0) {
            if (cuts > 0) out.emplace_back(o.begin() + start, o.end());
            break;
        }
        out.emplace_back(o.begin() + start, o.begin() + j + 1);  // points start..j, boundary repeated
        start = j;
        cuts++;
    }
    for (auto& s : out) {
        bool sorted = true;
        for (size_t k = 1; k < s.size(); k++)
            if (s[k].x < s[k - 1].x) sorted = false;
        if (!sorted) sortx(s);
    }
    return !out.empty();
}

// root of line a minus line b inside [p1,p2] (envelope.jl:199).  Inside one cell of the
// union grid both lines are linear, so the root is available in closed form; literal=1
// bisects the rounded difference to adjacent doubles like a bracketing fzero does.
double root_between(const Line& A, const Line& B, double p1, double p2, int literal, unsigned* status) {
    auto f = [&](double z) { return interp_y(A, z) - interp_y(B, z); };
    double f1 = f(p1), f2 = f(p2);
    if (f1 == 0.0) return p1;
    if (f2 == 0.0) return p2;
    if (!(f1 * f2 < 0.0)) {
        *status |= DCEGM_ST_NO_BRACKET;
        return p1;
    }
    if (!literal) {
        double r = p1 - f1 * (p2 - p1) / (f2 - f1);
        if (r < p1) r = p1;
        if (r > p2) r = p2;
        return r;
    }
    double a = p1, b = p2, fa = f1, fb = f2;
    for (int it = 0; it < 4000; it++) {
        double m = a + (b - a) / 2;
        if (!(m > a && m < b)) break;
        double fm = f(m);
        if (fm == 0.0) return m;
        if ((fm < 0) == (fa < 0)) {
            a = m;
            fa = fm;
        } else {
            b = m;
            fb = fm;
        }
    }
    return std::fabs(fa) <= std::fabs(fb) ? a : b;
}

// envelope.jl:143-258.  literal=1 follows the reference's own complexity (linear scans,
// bisection); literal=0 uses binary searches and the closed-form root.  Both must give the
// same envelope (tests/test_oracle_envelope.py).
Envelope upper_env(const std::vector<Line>& L, int literal, unsigned* status) {
    Envelope e;
    int S = (int)L.size();
    // xx = sort(unique(all x))                                       envelope.jl:154
    std::vector<double> xx;
    for (auto& l : L)
        for (auto& p : l) xx.push_back(p.x);
    std::sort(xx.begin(), xx.end());
    xx.erase(std::unique(xx.begin(), xx.end()), xx.end());
    int n = (int)xx.size();

    // yy = interp(L, xx, extrap=false); r_idx = linemax(yy); top    envelope.jl:158-164, line.jl:270-288
    std::vector<int> r_idx(n);
    std::vector<Pt> top(n);
    for (int j = 0; j < n; j++) {
        int iv = -1;
        double v = NEG_INF;
        Pt best{xx[j], NEG_INF};
        for (int s = 0; s < S; s++) {
            Pt p = interp_pt_noextrap(L[s], xx[j]);
            if (p.y > v) {
                v = p.y;
                iv = s;
                best = p;
            }
        }
        if (iv < 0) {
            *status |= DCEGM_ST_NONFINITE;
            iv = 0;
        }
        r_idx[j] = iv;
        top[j] = best;
    }

    auto is_knot = [&](const Line& l, double x) {
        if (literal) {
            for (auto& p : l)
                if (p.x == x) return true;
            return false;
        }
        int i = left_knot(l.data(), (int)l.size(), x);
        return l[i].x == x || l[i + 1].x == x;
    };
    auto in_domain = [&](const Line& l, double x) { return !(x < l.front().x || x > l.back().x); };

    e.env.push_back(top[0]);  // envelope.jl:170
    int k0 = r_idx[0];
    for (int j = 1; j < n; j++) {
        int k1 = r_idx[j];
        if (k0 != k1) {
            int ln1 = k0, ln2 = k1;
            double pt1 = xx[j - 1], pt2 = xx[j];
            // neither line extrapolated on [pt1,pt2]                 envelope.jl:183-188
            bool neither = in_domain(L[ln1], pt1) && in_domain(L[ln1], pt2) && in_domain(L[ln2], pt1) &&
                           in_domain(L[ln2], pt2);
            double fa = interp_y(L[ln1], pt1) - interp_y(L[ln2], pt1);
            double fb = interp_y(L[ln1], pt2) - interp_y(L[ln2], pt2);
            bool diffsign = fa * fb < 0;  // envelope.jl:191-192
            if (neither && diffsign) {
                int guard = 0;
                while (true) {
                    if (++guard > 4 * S + 16) {
                        *status |= DCEGM_ST_LOOP_LIMIT;
                        break;
                    }
                    unsigned st = 0;
                    double pt3 = root_between(L[ln1], L[ln2], pt1, pt2, literal, &st);  // envelope.jl:199
                    if (st) {
                        *status |= st;
                        break;
                    }
                    // all lines at pt3, first arg-max                envelope.jl:205-206
                    int ln3 = -1;
                    double y3 = NEG_INF;
                    for (int s = 0; s < S; s++) {
                        double y = interp_pt_noextrap(L[s], pt3).y;
                        if (ln3 < 0 || y > y3) {  // findmax: first maximal element
                            y3 = y;
                            ln3 = s;
                        }
                    }
                    if (ln3 == ln2 || ln3 == ln1) {
                        e.env.push_back(Pt{pt3, y3});     // envelope.jl:212
                        e.isects.push_back(Pt{pt3, y3});  // envelope.jl:214
                        if (ln2 == k1) break;              // envelope.jl:216-220
                        ln1 = ln2;                         // envelope.jl:225-228
                        pt1 = pt3;
                        ln2 = k1;
                        pt2 = xx[j];
                    } else {
                        ln2 = ln3;  // envelope.jl:236-237
                        pt2 = pt3;
                    }
                }
            }
        }
        // envelope.jl:248-252
        if (is_knot(L[k1], xx[j]) || j == n - 1) e.env.push_back(top[j]);
        k0 = k1;
    }
    return e;
}

// envelope.jl:64-71 — indices of d with no exactly equal Point in e.env
void removed_points(Envelope& e, const Line& d, int literal) {
    e.removed.clear();
    if (literal) {
        for (int k = 0; k < (int)d.size(); k++) {
            bool found = false;
            for (auto& q : e.env)
                if (q.x == d[k].x && q.y == d[k].y) {
                    found = true;
                    break;
                }
            if (!found) e.removed.push_back(k);
        }
        return;
    }
    // env is ascending in x when produced by upper_env; fall back to a scan if it is not
    bool sorted = true;
    for (size_t k = 1; k < e.env.size(); k++)
        if (e.env[k].x < e.env[k - 1].x) sorted = false;
    for (int k = 0; k < (int)d.size(); k++) {
        bool found = false;
        if (sorted) {
            auto lo = std::lower_bound(e.env.begin(), e.env.end(), d[k].x, [](const Pt& a, double x) { return a.x < x; });
            for (auto it = lo; it != e.env.end() && it->x == d[k].x; ++it)
                if (it->y == d[k].y) {
                    found = true;
                    break;
                }
        } else {
            for (auto& q : e.env)
                if (q.x == d[k].x && q.y == d[k].y) {
                    found = true;
                    break;
                }
        }
        if (!found) e.removed.push_back(k);
    }
}

// envelope.jl:268-285
Envelope secondary_envelope(const Line& L, int literal, unsigned* status) {
    std::vector<Line> A;
    bool ok = split_line(L, A);
    Envelope e;
    if (!ok) {
        *status |= DCEGM_ST_ALL_FALLING;
        e.env = L;
        e.n_segments = 0;
        return e;
    }
    if (A.size() > 1) {
        e = upper_env(A, literal, status);
    } else {
        e.env = A[0];
    }
    e.n_segments = (int)A.size();
    removed_points(e, L, literal);
    return e;
}

// dc_algo.jl:393-495.  vline/cline are modified like in the reference (bend-back prepend).
void do_secondary(Line& vline, Line& cline, bool working, double ev0, const Par& p, int literal, Envelope& vout,
                  Line& cout, unsigned* status) {
    double minx = std::numeric_limits<double>::max();  // line.jl:177-185
    for (auto& q : vline)
        if (q.x < minx) minx = q.x;
    int npre = 0;
    if (!(vline[0].x <= minx)) {
        // dc_algo.jl:401-418
        int k = p.na / 10;
        std::vector<double> x0;
        if (p.portable) {
            for (int i = 0; i < k; i++) x0.push_back(dc_linrange_at(minx, vline[0].x, k, i));
        } else {
            linrange(minx, vline[0].x, k, x0);
        }
        if (k > 0) x0.pop_back();
        std::vector<Pt> pv, pc;
        for (double x : x0) {
            double c0 = x;
            if (minx < 0) c0 = c0 + std::fabs(minx) + p.cfloor;
            double y0 = u_work(c0, working, p) + p.beta * ev0;
            pv.push_back(Pt{x, y0});
            pc.push_back(Pt{x, c0});
        }
        vline.insert(vline.begin(), pv.begin(), pv.end());
        cline.insert(cline.begin(), pc.begin(), pc.end());
        npre = (int)pv.size();
    }
    vout = secondary_envelope(vline, literal, status);
    vout.n_prepended = npre;
    cout = cline;

    if (!vout.removed.empty()) {
        const std::vector<int>& rm = vout.removed;
        std::vector<char> isrm(cout.size(), 0);
        for (int k : rm) isrm[k] = 1;
        std::vector<Pt> ins_left, ins_right;
        int nc = (int)cout.size();
        for (const Pt& I : vout.isects) {
            // from the left: largest surviving index with x < I.x       dc_algo.jl:442-462
            int jl = -1;
            for (int k = 0; k < nc; k++)
                if (cout[k].x < I.x && !isrm[k]) jl = k;
            if (jl >= 0 && jl < nc - 1) {
                Pt two[2] = {cout[jl], cout[jl + 1]};
                if (two[1].x < two[0].x) std::swap(two[0], two[1]);
                Pt t = interp_pt(two, 2, I.x);
                ins_left.push_back(Pt{t.x - 1e3 * 2.220446049250313e-16, t.y});
            }
            // from the right: smallest surviving index with x > I.x     dc_algo.jl:465-478
            int jr = -1;
            for (int k = nc - 1; k >= 0; k--)
                if (cout[k].x > I.x && !isrm[k]) jr = k;
            if (jr >= 0) {
                if (jr == 0) {
                    *status |= DCEGM_ST_INSERT_RANGE;
                } else {
                    Pt two[2] = {cout[jr - 1], cout[jr]};
                    if (two[1].x < two[0].x) std::swap(two[0], two[1]);
                    ins_right.push_back(interp_pt(two, 2, I.x));
                }
            }
        }
        // deleteat!(cout, rmidx)                                          dc_algo.jl:482
        Line kept;
        for (int k = 0; k < nc; k++)
            if (!isrm[k]) kept.push_back(cout[k]);
        cout.swap(kept);
        // the loop below runs over insert_left and indexes insert_right with it: a shorter insert_right is a
        // BoundsError, a longer one is silently ignored
        if (ins_right.size() < ins_left.size()) *status |= DCEGM_ST_MISALIGNED;
        if (ins_right.size() > ins_left.size()) *status |= DCEGM_ST_EXTRA_RIGHT;
        // dc_algo.jl:486-492
        for (size_t i = 0; i < ins_left.size(); i++) {
            int j = -1;
            for (int k = 0; k < (int)cout.size(); k++)
                if (cout[k].x > ins_left[i].x) {
                    j = k;
                    break;
                }
            if (j < 0 || i >= ins_right.size()) {
                *status |= DCEGM_ST_INSERT_RANGE;
                break;
            }
            cout.insert(cout.begin() + j, ins_left[i]);
            cout.insert(cout.begin() + j + 1, ins_right[i]);
        }
    }
}

// ---------------------------------------------------------------------------------------
// choice probabilities — dc_algo.jl:316-332 (two rows)
// ---------------------------------------------------------------------------------------
inline double ccp2(double v1, double v2, double lambda, int portable = 0) {
    double mx = v1 > v2 ? v1 : v2;  // maximum(x,dims=1)
    double e1 = m_exp((v1 - mx) / lambda, portable), e2 = m_exp((v2 - mx) / lambda, portable);
    return e1 / (e1 + e2);
}
inline double logsum2(double v1, double v2, double lambda, int portable = 0) {
    double mx = v1 > v2 ? v1 : v2;
    double e1 = m_exp((v1 - mx) / lambda, portable), e2 = m_exp((v2 - mx) / lambda, portable);
    return mx + lambda * m_log(e1 + e2, portable);
}

// dc_algo.jl:220-251.  itn is the 1-based period of the value line v (it+1 of the caller).
inline double vfun(int iid /*1-based*/, int itn, double c1, double m1, const Envelope& v, const Par& p) {
    if (itn == p.nT && p.nu > 0) return bequest_work(c1, false, p);
    bool mask = itn == p.nT ? true : (m1 < v.env[1].x);
    if (mask) return u_work(c1, iid == 1, p) + p.beta * v.vbound;
    return interp_y(v.env, m1);
}

// quadrature sum in the order fixed by quad_group (1: ascending iy; G>1: lane g sums
// iy = g, g+G, ... then a xor-butterfly over lanes G/2, ..., 1).  dc_algo.jl:58,74 leave
// the order to BLAS.
double quad_sum(const double* w, const double* f, int ny, int G) {
    if (G <= 1) {
        double s = w[0] * f[0];
        for (int iy = 1; iy < ny; iy++) s = s + w[iy] * f[iy];
        return s;
    }
    std::vector<double> lane(G, 0.0);
    for (int g = 0; g < G; g++) {
        bool first = true;
        double s = 0.0;
        for (int iy = g; iy < ny; iy += G) {
            if (first) {
                s = w[iy] * f[iy];
                first = false;
            } else
                s = s + w[iy] * f[iy];
        }
        lane[g] = s;
    }
    for (int off = G / 2; off >= 1; off >>= 1) {
        std::vector<double> nxt(G);
        for (int g = 0; g < G; g++) nxt[g] = lane[g] + lane[g ^ off];
        lane.swap(nxt);
    }
    return lane[0];
}

// ---------------------------------------------------------------------------------------
// output packing
// ---------------------------------------------------------------------------------------
struct OutView {
    dcegm_layout lay;
    double* lines;
    int32_t* lens;
    double* vbound;
    dcegm_env_meta* meta;
    double* isects;
    uint32_t* removed;
    double* dirty;
    int32_t* status;
};

inline size_t line_slot(const dcegm_layout& l, int m, int it, int iy, int id, int kind) {
    return ((((size_t)m * l.nT + it) * l.nstate + iy) * 2 + id) * 2 + kind;
}

void store_line(const OutView& o, int m, int it, int iy, int id, int kind, const Line& L, unsigned* status) {
    size_t s = line_slot(o.lay, m, it, iy, id, kind);
    int n = (int)L.size();
    if (n > o.lay.cap) {
        *status |= DCEGM_ST_LINE_OVERFLOW;
        n = o.lay.cap;
    }
    o.lens[s] = (int32_t)L.size();
    if (o.lines) {
        double* dst = o.lines + s * (size_t)o.lay.cap * 2;
        for (int k = 0; k < n; k++) {
            dst[2 * k] = L[k].x;
            dst[2 * k + 1] = L[k].y;
        }
    }
}
void store_dirty(const OutView& o, int m, int it, int iy, int id, int kind, const Line& L) {
    if (!o.dirty) return;
    size_t s = line_slot(o.lay, m, it, iy, id, kind);
    double* dst = o.dirty + s * (size_t)o.lay.na * 2;
    for (int k = 0; k < (int)L.size() && k < o.lay.na; k++) {
        dst[2 * k] = L[k].x;
        dst[2 * k + 1] = L[k].y;
    }
}
void store_env(const OutView& o, int m, int it, int iy, const Envelope& e, unsigned* status) {
    size_t s = ((size_t)m * o.lay.nT + it) * o.lay.nstate + iy;
    if (o.meta) {
        o.meta[s].n_segments = e.n_segments;
        o.meta[s].n_isects = (int)e.isects.size();
        o.meta[s].n_removed = (int)e.removed.size();
        o.meta[s].n_prepended = e.n_prepended;
    }
    if ((int)e.isects.size() > o.lay.isec_cap) *status |= DCEGM_ST_ISECT_OVERFLOW;
    if (o.isects) {
        double* dst = o.isects + s * (size_t)o.lay.isec_cap * 2;
        for (int k = 0; k < (int)e.isects.size() && k < o.lay.isec_cap; k++) {
            dst[2 * k] = e.isects[k].x;
            dst[2 * k + 1] = e.isects[k].y;
        }
    }
    if (o.removed) {
        uint32_t* dst = o.removed + s * (size_t)o.lay.rem_words;
        for (int k : e.removed)
            if (k / 32 < o.lay.rem_words) dst[k / 32] |= 1u << (k % 32);
    }
}

// tail of a period step — dc_algo.jl:94-101 / 194-201
// it: 1-based period of the step.  Lines of period 1 are never interpolated by the solve (dc_algo.jl:22-37 reads
// it+1 only): a defect there is informational (DCEGM_ST_NONMONO_FIRST), runf returns normally.
void finish_step(Envelope& v, Envelope& c, double a0, double ev0, int it, unsigned* status) {
    v.vbound = ev0;
    c.env.insert(c.env.begin(), Pt{a0, 0.0});
    v.env.insert(v.env.begin(), Pt{a0, ev0});
    sortx(c.env);
    // a line the previous period's interpolate() would reject
    *status |= knot_flags(c.env, it) | knot_flags(v.env, it);
    for (auto& q : c.env)
        if (!std::isfinite(q.x) || !std::isfinite(q.y)) *status |= DCEGM_ST_NONFINITE;
    for (auto& q : v.env)
        if (!std::isfinite(q.x) || !std::isfinite(q.y)) *status |= DCEGM_ST_NONFINITE;
}

// ---------------------------------------------------------------------------------------
// FModel — param.jl:189-223 and dc_EGM!(::FModel) dc_algo.jl:8-105
// ---------------------------------------------------------------------------------------
void solve_f_one(const dcegm_param& pin, int m, const OutView& o, int literal, int portable) {
    Par p = derive(pin, portable);
    unsigned status = 0;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const int G = p.quad_group > 0 ? p.quad_group : 1;
    // FModel(p)
    std::vector<double> nodes, ywgt, yvec(ny), avec;
    quadpoints(ny, 0, 1, nodes, ywgt);                                 // param.jl:194
    for (int i = 0; i < ny; i++) yvec[i] = norminv(nodes[i]) * p.sigma;  // param.jl:195-197
    linrange(p.a_low, p.a_high, na, avec);                             // param.jl:205

    std::vector<Envelope> c(2 * nT), v(2 * nT);  // [id + 2*it]
    auto at = [&](int id, int it) { return id + 2 * it; };
    const double NaN = std::numeric_limits<double>::quiet_NaN();

    std::vector<double> cm[2], vm[2], mu(ny), ls(ny), v2(ny);
    for (int it = nT; it >= 1; it--) {  // 1-based like the reference
        if (it == nT) {
            for (int id = 1; id <= 2; id++) {  // dc_algo.jl:14-20
                Envelope ec, evv;
                ec.env = {Pt{p.a_lowT, 0.0}, Pt{p.a_high, p.a_high}};
                if (p.nu > 0) {
                    for (int ia = 0; ia < na; ia++) evv.env.push_back(Pt{avec[ia], bequest_work(avec[ia], id == 1, p)});
                } else {
                    evv.env = {Pt{p.a_lowT, 0.0}, Pt{p.a_high, NaN}};
                }
                c[at(id - 1, it - 1)] = ec;
                v[at(id - 1, it - 1)] = evv;
            }
        } else {
            for (int id = 1; id <= 2; id++) {
                bool working = id == 1;
                const Envelope* cn[2] = {&c[at(0, it)], &c[at(1, it)]};  // period it+1
                const Envelope* vn[2] = {&v[at(0, it)], &v[at(1, it)]};
                if (vn[0]->env.size() < 2 || vn[1]->env.size() < 2) status |= DCEGM_ST_NONFINITE;
                std::vector<double> off(ny);
                for (int iy = 0; iy < ny; iy++) {  // param.jl:212-213
                    double inc = income(it + 1, p, yvec[iy]);
                    off[iy] = working ? inc : (p.delta * inc + (1 - p.delta) * p.pension);
                }
                bool beq = (it == nT - 1) && p.nu > 0;
                Line cline(na), vline(na);
                double ev0 = 0;
                for (int ia = 0; ia < na; ia++) {
                    for (int iy = 0; iy < ny; iy++) {
                        double m1 = avec[ia] * p.R + off[iy];
                        double c1[2], v1[2];
                        for (int iid = 1; iid <= 2; iid++) {  // dc_algo.jl:33-42
                            double cc = interp_y(cn[iid - 1]->env, m1);
                            if (cc < p.cfloor) cc = p.cfloor;  // floory!, line.jl:131-139
                            c1[iid - 1] = cc;
                            v1[iid - 1] = vfun(iid, it + 1, cc, m1, *vn[iid - 1], p);
                        }
                        double pwork = working ? ccp2(v1[0], v1[1], p.lambda, p.portable) : p.delta;  // dc_algo.jl:44
                        if (beq) {                                                          // dc_algo.jl:46-48
                            c1[0] = c1[0] + p.bbar;
                            c1[1] = c1[1] + p.bbar;
                        }
                        mu[iy] = pwork * up(c1[0], p) + (1.0 - pwork) * up(c1[1], p);  // dc_algo.jl:49,53
                        ls[iy] = logsum2(v1[0], v1[1], p.lambda, p.portable);
                        v2[iy] = v1[1];
                    }
                    double emu = quad_sum(ywgt.data(), mu.data(), ny, G);
                    double rhs = beq ? (p.beta * p.R * p.nu) * emu : (p.beta * p.R) * emu;  // dc_algo.jl:50,58
                    double c0 = iup(rhs, p);                                                 // dc_algo.jl:62
                    double ev;
                    if (working)
                        ev = quad_sum(ywgt.data(), ls.data(), ny, G);  // dc_algo.jl:74
                    else
                        ev = (1 - p.delta) * quad_sum(ywgt.data(), v2.data(), ny, G) +
                             p.delta * quad_sum(ywgt.data(), ls.data(), ny, G);  // dc_algo.jl:76
                    if (ia == 0) ev0 = ev;
                    cline[ia] = Pt{avec[ia] + c0, c0};                                  // dc_algo.jl:65
                    vline[ia] = Pt{avec[ia] + c0, u_work(c0, id == 1, p) + p.beta * ev};  // dc_algo.jl:78
                }
                store_dirty(o, m, it - 1, 0, id - 1, DCEGM_KIND_C, cline);
                store_dirty(o, m, it - 1, 0, id - 1, DCEGM_KIND_V, vline);
                Envelope evn, ecn;
                if (id == 1) {  // dc_algo.jl:87-88
                    Line cout;
                    do_secondary(vline, cline, working, ev0, p, literal, evn, cout, &status);
                    ecn.env = cout;
                    store_env(o, m, it - 1, 0, evn, &status);
                } else {
                    evn.env = vline;
                    ecn.env = cline;
                }
                finish_step(evn, ecn, avec[0], ev0, it, &status);
                c[at(id - 1, it - 1)] = ecn;
                v[at(id - 1, it - 1)] = evn;
            }
        }
        for (int id = 0; id < 2; id++) {
            store_line(o, m, it - 1, 0, id, DCEGM_KIND_C, c[at(id, it - 1)].env, &status);
            store_line(o, m, it - 1, 0, id, DCEGM_KIND_V, v[at(id, it - 1)].env, &status);
            o.vbound[(((size_t)m * nT + (it - 1)) * 1 + 0) * 2 + id] = v[at(id, it - 1)].vbound;
        }
    }
    o.status[m] = (int32_t)status;
}

// ---------------------------------------------------------------------------------------
// GModel — param.jl:261-310 and dc_EGM!(::GModel) dc_algo.jl:112-206
// ---------------------------------------------------------------------------------------
void solve_g_one(const dcegm_param& pin, int m, const OutView& o, int literal, int portable) {
    Par p = derive(pin, portable);
    unsigned status = 0;
    const int na = p.na, ny = p.ny, nT = p.nT;
    std::vector<double> yvec, ywgt /* row-major [iy][jy] */, avec;
    if (p.rho == 0) {  // param.jl:278-281
        std::vector<double> nodes, w;
        gauss_hermite(ny, nodes, w);
        yvec.resize(ny);
        ywgt.resize((size_t)ny * ny);
        const double pi = 3.141592653589793;
        for (int i = 0; i < ny; i++) yvec[i] = std::sqrt(2.0) * p.sigma * nodes[i];
        for (int iy = 0; iy < ny; iy++)
            for (int jy = 0; jy < ny; jy++) ywgt[(size_t)iy * ny + jy] = w[jy] * std::pow(pi, -0.5);
    } else {
        rouwenhorst(p.rho, 0.0, p.sigma, ny, yvec, ywgt);  // param.jl:284
    }
    scale_grid(p.a_low, p.a_high, na, 2, avec);  // param.jl:288

    std::vector<Envelope> c((size_t)2 * ny * nT), v((size_t)2 * ny * nT);
    auto at = [&](int id, int iy, int it) { return (size_t)id + 2 * ((size_t)iy + (size_t)ny * it); };
    const double NaN = std::numeric_limits<double>::quiet_NaN();

    std::vector<double> vmat[2], cmat[2];
    for (int k = 0; k < 2; k++) {
        vmat[k].resize(na);
        cmat[k].resize(na);
    }
    for (int it = nT; it >= 1; it--) {
        bool is_bequest = (it == nT - 1) && (p.nu > 0);
        if (it == nT) {  // dc_algo.jl:122-131
            for (int id = 1; id <= 2; id++)
                for (int iy = 0; iy < ny; iy++) {
                    Envelope ec, evv;
                    ec.env = {Pt{p.a_lowT, 0.0}, Pt{p.a_high, p.a_high}};
                    if (p.nu > 0) {
                        for (int ia = 0; ia < na; ia++)
                            evv.env.push_back(Pt{avec[ia], bequest_work(avec[ia], id == 1, p) - (id > 1 ? 10.0 : 0.0)});
                    } else {
                        evv.env = {Pt{p.a_lowT, 0.0}, Pt{p.a_high, NaN}};
                    }
                    c[at(id - 1, iy, it - 1)] = ec;
                    v[at(id - 1, iy, it - 1)] = evv;
                }
        } else {
            for (int iy = 0; iy < ny; iy++) {
                for (int k = 0; k < 2; k++) {
                    std::fill(vmat[k].begin(), vmat[k].end(), 0.0);
                    std::fill(cmat[k].begin(), cmat[k].end(), 0.0);
                }
                for (int jy = 0; jy < ny; jy++) {  // dc_algo.jl:142-157
                    double pr = ywgt[(size_t)iy * ny + jy];
                    for (int iid = 1; iid <= 2; iid++) {
                        const Envelope& cn = c[at(iid - 1, jy, it)];
                        const Envelope& vn = v[at(iid - 1, jy, it)];
                        double inc = income(it + 1, p, yvec[jy]);
                        for (int ia = 0; ia < na; ia++) {
                            // param.jl:301
                            double m1 = (avec[ia] * p.R) + (iid == 1 ? inc : 0.0) + (iid == 2 ? p.pension : 0.0);
                            double cc = interp_y(cn.env, m1);
                            if (cc < p.cfloor) cc = p.cfloor;
                            vmat[iid - 1][ia] += pr * vfun(iid, it + 1, cc, m1, vn, p);
                            if (is_bequest)
                                cmat[iid - 1][ia] += pr * up(cc + p.bbar, p);
                            else
                                cmat[iid - 1][ia] += pr * up(cc, p);
                        }
                    }
                }
                for (int id = 1; id <= 2; id++) {  // dc_algo.jl:159-202
                    bool working = id == 1;
                    Line cline(na), vline(na);
                    double ev0 = 0;
                    for (int ia = 0; ia < na; ia++) {
                        double pwork = working ? ccp2(vmat[0][ia], vmat[1][ia], p.lambda, p.portable) : 0.0;
                        double mu1 = pwork * cmat[0][ia] + (1.0 - pwork) * cmat[1][ia];
                        double rhs = is_bequest ? (p.beta * p.R * p.nu) * mu1 : (p.beta * p.R) * mu1;
                        double c0 = iup(rhs, p);
                        double ev = working ? logsum2(vmat[0][ia], vmat[1][ia], p.lambda, p.portable) : vmat[1][ia];
                        if (ia == 0) ev0 = ev;
                        cline[ia] = Pt{avec[ia] + c0, c0};
                        vline[ia] = Pt{avec[ia] + c0, u_work(c0, id == 1, p) + p.beta * ev};
                    }
                    store_dirty(o, m, it - 1, iy, id - 1, DCEGM_KIND_C, cline);
                    store_dirty(o, m, it - 1, iy, id - 1, DCEGM_KIND_V, vline);
                    Envelope evn, ecn;
                    if (id == 1) {
                        Line cout;
                        do_secondary(vline, cline, working, ev0, p, literal, evn, cout, &status);
                        ecn.env = cout;
                        store_env(o, m, it - 1, iy, evn, &status);
                    } else {
                        evn.env = vline;
                        ecn.env = cline;
                    }
                    finish_step(evn, ecn, avec[0], ev0, it, &status);
                    c[at(id - 1, iy, it - 1)] = ecn;
                    v[at(id - 1, iy, it - 1)] = evn;
                }
            }
        }
        for (int iy = 0; iy < ny; iy++)
            for (int id = 0; id < 2; id++) {
                store_line(o, m, it - 1, iy, id, DCEGM_KIND_C, c[at(id, iy, it - 1)].env, &status);
                store_line(o, m, it - 1, iy, id, DCEGM_KIND_V, v[at(id, iy, it - 1)].env, &status);
                o.vbound[(((size_t)m * nT + (it - 1)) * ny + iy) * 2 + id] = v[at(id, iy, it - 1)].vbound;
            }
    }
    o.status[m] = (int32_t)status;
}

// ---------------------------------------------------------------------------------------
// minimal_EGM — minimal_EGM.jl:3-43 (data only; the reference returns a plot)
// ---------------------------------------------------------------------------------------
// minimal_EGM (minimal_EGM.jl:3-43), minimal_EGM_borrow (:49-89), minimal_EGM_bequest (:92-159).
// No reference test covers them: parity unpinned.
void solve_minimal_one(const dcegm_param& pin, int m, const OutView& o, int model, int portable) {
    Par p = derive(pin, portable);
    unsigned status = 0;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const bool borrow = model == DCEGM_MODEL_MINIMAL_BORROW, beq = model == DCEGM_MODEL_MINIMAL_BEQUEST;
    std::vector<double> nodes, w, yvec(ny), ywgt(ny), avec;
    gauss_hermite(ny, nodes, w);  // minimal_EGM.jl:4-6
    const double pi = 3.141592653589793;
    for (int i = 0; i < ny; i++) {
        yvec[i] = std::sqrt(2.0) * p.sigma * nodes[i];
        ywgt[i] = w[i] * std::pow(pi, -0.5);
    }
    const double m_lo = borrow ? p.a_low : 0.0;
    if (beq)
        scale_grid(0.0, p.a_high, na, 1, avec);  // minimal_EGM.jl:97
    else
        linrange(m_lo, p.a_high, na, avec);      // minimal_EGM.jl:7, 53
    Line next, nextv;
    if (beq) {  // minimal_EGM.jl:102-104
        next.resize(na);
        nextv.resize(na);
        for (int k = 0; k < na; k++) {
            next[k] = Pt{avec[k], avec[k]};
            nextv[k] = Pt{avec[k], p.nu > 0 ? p.nu * u_plain(avec[k] + p.bbar, p) : 0.0};
        }
        store_line(o, m, nT - 1, 0, 0, DCEGM_KIND_V, nextv, &status);
    } else {
        next = {Pt{m_lo, 0.0}, Pt{p.a_high, p.a_high}};  // minimal_EGM.jl:10-11, 56-57
    }
    store_line(o, m, nT - 1, 0, 0, DCEGM_KIND_C, next, &status);
    std::vector<double> ey(ny), inv(ny), v1(ny);
    for (int i = 0; i < ny; i++) ey[i] = 0.0 + std::exp(yvec[i]);
    const int G = p.quad_group > 0 ? p.quad_group : 1;
    for (int it = nT - 1; it >= 1; it--) {
        const bool beq_period = beq && it == nT - 1 && p.nu > 0;  // minimal_EGM.jl:124-133
        Line cur(na + 1), curv(na + 1);
        cur[0] = Pt{m_lo, 0.0};  // minimal_EGM.jl:36-37, 82-83
        double ev0 = 0.0;
        for (int ia = 0; ia < na; ia++) {
            for (int iy = 0; iy < ny; iy++) {
                double w1 = ey[iy] + p.R * avec[ia];  // minimal_EGM.jl:21
                double c1 = interp_y(next, w1);        // minimal_EGM.jl:25
                if (c1 < 0) c1 = p.cfloor;             // minimal_EGM.jl:26
                inv[iy] = beq_period ? 1.0 / (c1 + p.bbar) : 1.0 / c1;
                if (beq) v1[iy] = interp_y(nextv, w1);  // minimal_EGM.jl:150
            }
            double emu = quad_sum(ywgt.data(), inv.data(), ny, G);            // minimal_EGM.jl:27
            double rhs = (beq_period ? (p.beta * p.R * p.nu) : (p.beta * p.R)) * emu;  // minimal_EGM.jl:28, 132
            double c0 = 1.0 / rhs;                                            // minimal_EGM.jl:29
            cur[ia + 1] = Pt{avec[ia] + c0, c0};                              // minimal_EGM.jl:30
            if (beq) {
                double ev = quad_sum(ywgt.data(), v1.data(), ny, G);          // minimal_EGM.jl:151
                curv[ia + 1] = Pt{avec[ia] + c0, u_plain(c0, p) + p.beta * ev};  // minimal_EGM.jl:152
                if (ia == 0) ev0 = ev;
            }
        }
        status |= knot_flags(cur, it);
        store_line(o, m, it - 1, 0, 0, DCEGM_KIND_C, cur, &status);
        if (beq) {
            curv[0] = Pt{m_lo, ev0};
            store_line(o, m, it - 1, 0, 0, DCEGM_KIND_V, curv, &status);
            nextv.swap(curv);
        }
        next.swap(cur);
    }
    o.status[m] = (int32_t)status;
}


// ---------------------------------------------------------------------------------------
// forward simulation — simulate.jl:61-130 (FModel) and :133-203 (GModel)
// ---------------------------------------------------------------------------------------
// Parity unpinned: the reference has no test of sim(), and its shocks come from Julia's global RNG.
// Shocks here are either given (the raw rand() draws, column-major [agent + nsims*it]) or the
// counter-based draws of dc_rng.h (the CUDA library's definition, shared header).  Per-period sums
// follow the CUDA kernel's fixed order: tiles of 1024 agents, thread t of 256 owns agents
// tile0 + a*256 + t (a = 0..3, ascending), 32-lane xor butterfly, warps ascending, tiles ascending.
struct SimLines {
    const dcegm_layout* lay;
    const double* lines;
    const int32_t* lens;
    const double* vbound;
    int m;
    const Pt* line(int it, int st, int id, int kind, int* n) const {
        size_t slot = ((((size_t)m * lay->nT + it) * lay->nstate + st) * 2 + id) * 2 + kind;
        *n = lens[slot] < lay->cap ? lens[slot] : lay->cap;
        return reinterpret_cast<const Pt*>(lines + slot * (size_t)lay->cap * 2);
    }
    double vb(int it, int st, int id) const { return vbound[(((size_t)m * lay->nT + it) * lay->nstate + st) * 2 + id]; }
};

// vfun (dc_algo.jl:220-251) on a flat line; it0 is the 0-based period of the line
double sim_vfun(bool worker_line, bool terminal, double c1, double m1, const Pt* v, int n, double vb, const Par& p) {
    if (terminal && p.nu > 0) return bequest_work(c1, false, p);
    bool mask = terminal || (m1 < v[1].x);
    if (mask) return u_work(c1, worker_line, p) + p.beta * vb;
    return interp_pt(v, n, m1).y;
}

double sim_income(int it1, const Par& p, double shock) {  // utils.jl:243-250
    int age = it1 + 19;
    if (it1 > p.retage) return p.pension;
    return m_exp(p.inc0 + p.inc1 * (double)age - p.inc2 * (double)(age * age) + shock, p.portable);
}

void simulate_one(int model, const dcegm_param& pin, const SimLines& S, int portable, const dcegm_sim_shocks* sh,
                  double* moments /*[nT][8]*/, int32_t* ret_hist /*[nT+1]*/, int32_t* status_out, double* pan_d,
                  int32_t* pan_i, int32_t* pan_ret) {
    Par p = derive(pin, portable);
    const int nT = p.nT, ny = p.ny;
    const int nsims = sh ? sh->nsims : p.nsims;
    const bool G = model == DCEGM_MODEL_G;
    const int TPB = 256, APT = 4, TILE = TPB * APT, NM = DCEGM_SIM_NMOM;
    unsigned status = 0;
    std::vector<double> gg, inctab;
    if (G) {
        std::vector<double> yvec, ywgt;
        if (p.rho == 0) {  // param.jl:278-281
            std::vector<double> nodes, w;
            gauss_hermite(ny, nodes, w);
            yvec.resize(ny);
            ywgt.resize((size_t)ny * ny);
            const double pi = 3.141592653589793;
            for (int i = 0; i < ny; i++) yvec[i] = std::sqrt(2.0) * p.sigma * nodes[i];
            for (int iy = 0; iy < ny; iy++)
                for (int jy = 0; jy < ny; jy++) ywgt[(size_t)iy * ny + jy] = w[jy] * std::pow(pi, -0.5);
        } else {
            rouwenhorst(p.rho, 0.0, p.sigma, ny, yvec, ywgt);
        }
        gg.resize((size_t)ny * ny);  // cumsum(m.ywgt, dims=2), simulate.jl:145
        for (int r = 0; r < ny; r++) {
            double acc = 0.0;
            for (int j = 0; j < ny; j++) {
                acc = j == 0 ? ywgt[(size_t)r * ny + j] : acc + ywgt[(size_t)r * ny + j];
                gg[(size_t)r * ny + j] = acc;
            }
        }
        inctab.resize((size_t)nT * ny);
        for (int it = 0; it < nT; it++)
            for (int iy = 0; iy < ny; iy++) inctab[(size_t)it * ny + iy] = sim_income(it + 1, p, yvec[iy]);
    }
    std::vector<double> mom((size_t)nT * NM, 0.0);
    std::vector<int32_t> hist(nT + 1, 0);
    const size_t npan = (size_t)nsims * nT;
    const double NaN = std::numeric_limits<double>::quiet_NaN();
    std::vector<double> part((size_t)nT * NM * TPB);  // per-thread partial sums of the current tile
    for (int tile0 = 0; tile0 < nsims; tile0 += TILE) {
        std::fill(part.begin(), part.end(), 0.0);
        for (int t = 0; t < TPB; t++)
            for (int a = 0; a < APT; a++) {
                const int i = tile0 + a * TPB + t;
                if (i >= nsims) continue;
                double w1 = 0.0;
                int ys = 0, ret = 0;
                bool working = true, prev_worker = true;
                for (int it = 0; it < nT; it++) {
                    const size_t at = (size_t)i + (size_t)nsims * it;
                    const bool wk = working;
                    double w0, inc;
                    if (G) {
                        if (it == 0) {
                            int y0 = sh ? sh->ystate0[i] - 1
                                        : (int)(dc_sim_uniform(p.rseed, (uint32_t)i, 0u, DC_STREAM_Y0) * (double)ny);  // rand(1:ny), simulate.jl:141
                            ys = y0 < 0 ? 0 : (y0 > ny - 1 ? ny - 1 : y0);
                        } else {  // simulate.jl:148
                            const double uw = sh ? sh->wshocks[at] : dc_sim_uniform(p.rseed, (uint32_t)i, (uint32_t)it, DC_STREAM_W);
                            const double* row = &gg[(size_t)ys * ny];
                            int j = 0;
                            while (j < ny && !(row[j] >= uw)) j++;
                            if (j >= ny) {
                                status |= DCEGM_ST_YSTATE_RANGE;
                                j = ny - 1;
                            }
                            ys = j;
                        }
                    }
                    if (it == 0) {  // simulate.jl:82-86, 160-164
                        const double u0 = sh ? sh->w0shocks[i] : dc_sim_uniform(p.rseed, (uint32_t)i, 0u, DC_STREAM_W0);
                        w0 = p.initw0 + u0 * (p.initw1 - p.initw0);
                        inc = G ? inctab[ys] : NaN;
                    } else {
                        if (prev_worker && !wk) ret = it + 1;  // simulate.jl:95
                        if (!wk) {
                            inc = p.pension;
                        } else if (G) {
                            inc = inctab[(size_t)it * ny + ys];
                        } else {
                            const double uw = sh ? sh->wshocks[at] : dc_sim_uniform(p.rseed, (uint32_t)i, (uint32_t)it, DC_STREAM_W);
                            inc = sim_income(it + 1, p, dc_norm_quantile(uw) * p.sigma);  // simulate.jl:73,99
                        }
                        w0 = w1 * p.R + inc;  // simulate.jl:103
                    }
                    const int st = G ? ys : 0, id = wk ? 0 : 1;
                    int n;
                    const Pt* cl = S.line(it, st, id, 0, &n);
                    const double cons = interp_pt(cl, n, w0).y;  // simulate.jl:106-110
                    const double util = u_work(cons, wk, p);
                    double w1n = w0 - cons;
                    if (!G) w1n = w1n > p.a_high ? p.a_high : (w1n < p.a_low ? p.a_low : w1n);  // clamp!, simulate.jl:109
                    const bool negc = cons < 0.0;
                    if (negc && !G) status |= DCEGM_ST_NEG_CONS;
                    const bool term = it == nT - 1;
                    int n1, n2;
                    const Pt* l1 = S.line(it, st, 0, 1, &n1);
                    const Pt* l2 = S.line(it, st, 1, 1, &n2);
                    const double v1 = sim_vfun(true, term, cons, w0, l1, n1, S.vb(it, st, 0), p);
                    const double v2 = sim_vfun(false, term, cons, w0, l2, n2, S.vb(it, st, 1), p);
                    const double pw = wk ? ccp2(v1, v2, p.lambda, p.portable) : 0.0;  // simulate.jl:121
                    const double ud = sh ? sh->dshocks[at] : dc_sim_uniform(p.rseed, (uint32_t)i, (uint32_t)it, DC_STREAM_D);
                    bool next = pw > ud;  // simulate.jl:124
                    if (!G && !wk) {      // simulate.jl:127
                        const double ur = sh ? (sh->rshocks ? sh->rshocks[at] : 1.0) : dc_sim_uniform(p.rseed, (uint32_t)i, (uint32_t)it, DC_STREAM_R);
                        next = p.delta > ur;
                    }
                    prev_worker = wk;
                    working = next;
                    w1 = w1n;
                    double* ps = &part[((size_t)it * NM) * TPB + t];
                    const double vals[8] = {w0, w1n, cons, util, inc, pw, wk ? 1.0 : 0.0, negc ? 1.0 : 0.0};
                    for (int k = 0; k < NM; k++) ps[(size_t)k * TPB] += vals[k];
                    if (pan_d) {
                        for (int k = 0; k < 6; k++) pan_d[(size_t)k * npan + at] = vals[k];
                        pan_i[at] = G ? ys + 1 : 0;
                        pan_i[npan + at] = wk ? 1 : 2;
                    }
                }
                hist[ret]++;
                if (pan_ret) pan_ret[i] = ret;
            }
        for (int it = 0; it < nT; it++)
            for (int k = 0; k < NM; k++) {
                const double* v = &part[((size_t)it * NM + k) * TPB];
                double tsum = 0.0;
                for (int w = 0; w < TPB / 32; w++) {
                    double lane[32], nxt[32];
                    for (int l = 0; l < 32; l++) lane[l] = v[w * 32 + l];
                    for (int o = 16; o > 0; o >>= 1) {
                        for (int l = 0; l < 32; l++) nxt[l] = lane[l] + lane[l ^ o];
                        for (int l = 0; l < 32; l++) lane[l] = nxt[l];
                    }
                    tsum = w == 0 ? lane[0] : tsum + lane[0];
                }
                mom[(size_t)it * NM + k] = tile0 == 0 ? tsum : mom[(size_t)it * NM + k] + tsum;
            }
    }
    const double dn = (double)nsims;
    for (size_t k = 0; k < mom.size(); k++) moments[k] = mom[k] / dn;
    for (int k = 0; k <= nT; k++) ret_hist[k] = hist[k];
    *status_out = (int32_t)status;
}

}  // namespace

// =========================================================================================
// exported C interface of the oracle (ctypes: oracle/oracle.py)
// =========================================================================================
extern "C" {
void oracle_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out) {
    dc_philox4x32_10(c0, c1, c2, c3, k0, k1, out);
}
double oracle_sim_uniform(int64_t seed, uint32_t agent, uint32_t period, uint32_t stream) {
    return dc_sim_uniform(seed, agent, period, stream);
}
double oracle_norm_quantile(double p) { return dc_norm_quantile(p); }

// forward simulation of models solved in the shared flat layout (lines/lens/vbound as written by
// oracle_solve or fetched from the CUDA library).  panel_model >= 0 also fills that model's panel.
int oracle_simulate(int model, const dcegm_param* params, int nmodels, int portable, const dcegm_layout* lay,
                    const double* lines, const int32_t* lens, const double* vbound, const dcegm_sim_shocks* shocks,
                    int nthreads, double* moments, int32_t* ret_hist, int32_t* sim_status, int panel_model, double* pan_d,
                    int32_t* pan_i, int32_t* pan_ret) {
    if (model != DCEGM_MODEL_F && model != DCEGM_MODEL_G) return -1;
    const int nT = lay->nT;
    std::atomic<int> next(0);
    auto work = [&]() {
        for (;;) {
            int m = next.fetch_add(1);
            if (m >= nmodels) break;
            SimLines S{lay, lines, lens, vbound, m};
            const bool pan = m == panel_model;
            simulate_one(model, params[m], S, portable, shocks, moments + (size_t)m * nT * DCEGM_SIM_NMOM,
                         ret_hist + (size_t)m * (nT + 1), sim_status + m, pan ? pan_d : nullptr, pan ? pan_i : nullptr,
                         pan ? pan_ret : nullptr);
        }
    };
    if (nthreads <= 1)
        work();
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; t++) pool.emplace_back(work);
        for (auto& th : pool) th.join();
    }
    return 0;
}


// geometry shared with the CUDA library (same formula as dcegm.jl_b200/csrc): the tests
// compare flat buffers slot by slot.
void oracle_layout(int model, int nmodels, int na, int ny, int nT, int keep_dirty, dcegm_layout* l) {
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
    l->seg_cap = 0;
    l->has_dirty = keep_dirty;
    l->has_lines = 1;
}

// Solve a batch on `nthreads` host threads (std::thread).  Buffers follow include/dcegm.h.
int oracle_solve(int model, const dcegm_param* params, int nmodels, int keep_dirty, int literal, int portable,
                 int nthreads,
                 double* lines, int32_t* lens, double* vbound, dcegm_env_meta* meta, double* isects,
                 uint32_t* removed, double* dirty, int32_t* status) {
    if (nmodels <= 0) return 0;
    OutView o;
    oracle_layout(model, nmodels, params[0].na, params[0].ny, params[0].nT, keep_dirty, &o.lay);
    o.lines = lines;
    o.lens = lens;
    o.vbound = vbound;
    o.meta = meta;
    o.isects = isects;
    o.removed = removed;
    o.dirty = keep_dirty ? dirty : nullptr;
    o.status = status;
    for (int m = 0; m < nmodels; m++)
        if (params[m].na != params[0].na || params[m].ny != params[0].ny || params[m].nT != params[0].nT) return -1;
    // one model per task, dynamic schedule over `nthreads` host threads
    std::atomic<int> next(0);
    auto worker = [&]() {
        for (;;) {
            int m = next.fetch_add(1);
            if (m >= nmodels) break;
            if (model == DCEGM_MODEL_F)
                solve_f_one(params[m], m, o, literal, portable);
            else if (model == DCEGM_MODEL_G)
                solve_g_one(params[m], m, o, literal, portable);
            else
                solve_minimal_one(params[m], m, o, model, portable);
        }
    };
    int nt = nthreads > 0 ? nthreads : 1;
    if (nt > nmodels) nt = nmodels;
    if (nt <= 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    return 0;
}

// ---- unit-level hooks used by the known-answer tests -------------------------------------
// secondary_envelope on one line.  Outputs: env (cap_env points), isects, removed indices.
int oracle_secondary_envelope(const double* x, const double* y, int n, int literal, double* env_xy, int cap_env,
                              int* n_env, double* isec_xy, int cap_isec, int* n_isec, int* removed, int cap_rem,
                              int* n_rem, int* n_seg) {
    Line L(n);
    for (int i = 0; i < n; i++) L[i] = Pt{x[i], y[i]};
    unsigned st = 0;
    Envelope e = secondary_envelope(L, literal, &st);
    *n_env = (int)e.env.size();
    *n_isec = (int)e.isects.size();
    *n_rem = (int)e.removed.size();
    *n_seg = e.n_segments;
    for (int i = 0; i < *n_env && i < cap_env; i++) {
        env_xy[2 * i] = e.env[i].x;
        env_xy[2 * i + 1] = e.env[i].y;
    }
    for (int i = 0; i < *n_isec && i < cap_isec; i++) {
        isec_xy[2 * i] = e.isects[i].x;
        isec_xy[2 * i + 1] = e.isects[i].y;
    }
    for (int i = 0; i < *n_rem && i < cap_rem; i++) removed[i] = e.removed[i];
    return (int)st;
}

// upper_env on explicit sections: seg_off[nseg+1] offsets into x/y.
int oracle_upper_env(const double* x, const double* y, const int* seg_off, int nseg, int literal, double* env_xy,
                     int cap_env, int* n_env, double* isec_xy, int cap_isec, int* n_isec) {
    std::vector<Line> A(nseg);
    for (int s = 0; s < nseg; s++)
        for (int i = seg_off[s]; i < seg_off[s + 1]; i++) A[s].push_back(Pt{x[i], y[i]});
    unsigned st = 0;
    Envelope e = upper_env(A, literal, &st);
    *n_env = (int)e.env.size();
    *n_isec = (int)e.isects.size();
    for (int i = 0; i < *n_env && i < cap_env; i++) {
        env_xy[2 * i] = e.env[i].x;
        env_xy[2 * i + 1] = e.env[i].y;
    }
    for (int i = 0; i < *n_isec && i < cap_isec; i++) {
        isec_xy[2 * i] = e.isects[i].x;
        isec_xy[2 * i + 1] = e.isects[i].y;
    }
    return (int)st;
}

// splitLine: returns number of sections; seg_len[s], seg_xmin[s], seg_xmax[s]
int oracle_split_line(const double* x, const double* y, int n, int* seg_len, double* seg_xmin, double* seg_xmax,
                      int cap) {
    Line L(n);
    for (int i = 0; i < n; i++) L[i] = Pt{x[i], y[i]};
    std::vector<Line> A;
    split_line(L, A);
    for (int s = 0; s < (int)A.size() && s < cap; s++) {
        seg_len[s] = (int)A[s].size();
        seg_xmin[s] = A[s].front().x;
        seg_xmax[s] = A[s].back().x;
    }
    return (int)A.size();
}

void oracle_interp(const double* x, const double* y, int n, const double* q, int nq, double* out_x, double* out_y) {
    Line L(n);
    for (int i = 0; i < n; i++) L[i] = Pt{x[i], y[i]};
    for (int i = 0; i < nq; i++) {
        Pt r = interp_pt(L.data(), n, q[i]);
        out_x[i] = r.x;
        out_y[i] = r.y;
    }
}

// ccp / logsum on a 2 x n matrix given as two rows — dc_algo.jl:316-332
void oracle_ccp(const double* v1, const double* v2, int n, double lambda, double* out) {
    for (int i = 0; i < n; i++) out[i] = ccp2(v1[i], v2[i], lambda);
}
void oracle_logsum(const double* v1, const double* v2, int n, double lambda, double* out) {
    for (int i = 0; i < n; i++) out[i] = logsum2(v1[i], v2[i], lambda);
}

void oracle_quadpoints(int n, double lo, double hi, double* x, double* w) {
    std::vector<double> xx, ww;
    quadpoints(n, lo, hi, xx, ww);
    std::copy(xx.begin(), xx.end(), x);
    std::copy(ww.begin(), ww.end(), w);
}
double oracle_norminv(double p) { return norminv(p); }
void oracle_gausshermite(int n, double* x, double* w) {
    std::vector<double> xx, ww;
    gauss_hermite(n, xx, ww);
    std::copy(xx.begin(), xx.end(), x);
    std::copy(ww.begin(), ww.end(), w);
}
void oracle_rouwenhorst(double rho, double mu, double sigma, int n, double* z, double* P) {
    std::vector<double> zz, PP;
    rouwenhorst(rho, mu, sigma, n, zz, PP);
    std::copy(zz.begin(), zz.end(), z);
    std::copy(PP.begin(), PP.end(), P);
}
void oracle_scalegrid(double lb, double ub, int n, int logorder, double* out) {
    std::vector<double> g;
    scale_grid(lb, ub, n, logorder, g);
    std::copy(g.begin(), g.end(), out);
}
void oracle_linrange(double lo, double hi, int n, double* out) {
    std::vector<double> g;
    linrange(lo, hi, n, g);
    std::copy(g.begin(), g.end(), out);
}
// the shared reproducible math, exported for accuracy tests against mpmath
void oracle_dc_pow(const double* x, const double* y, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = dc_pow(x[i], y[i]);
}
void oracle_dc_exp(const double* x, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = dc_exp(x[i]);
}
void oracle_dc_log(const double* x, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = dc_log(x[i]);
}
double oracle_dc_linrange_at(double lo, double hi, int n, int i) { return dc_linrange_at(lo, hi, n, i); }
double oracle_u(double x, int working, const dcegm_param* p) { return u_work(x, working != 0, derive(*p)); }
double oracle_up(double x, const dcegm_param* p) { return up(x, derive(*p)); }
double oracle_iup(double x, const dcegm_param* p) { return iup(x, derive(*p)); }
double oracle_income(int it, double shock, const dcegm_param* p) { return income(it, derive(*p), shock); }

}  // extern "C"
extern "C" void oracle_dc_log_parts(const double* x, int n, double* hi, double* lo) {
    for (int i = 0; i < n; i++) dc_log_parts(x[i], &hi[i], &lo[i]);
}
extern "C" void oracle_dc_exp_parts(const double* h, const double* l, int n, double* out) {
    for (int i = 0; i < n; i++) out[i] = dc_exp_parts(h[i], l[i]);
}
