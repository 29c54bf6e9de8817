// host_setup.cpp — see host_setup.hpp.  Plain C++17, no CUDA.
#include "host_setup.hpp"

#include <algorithm>
#include <cmath>
#include <limits>
#include <map>
#include <thread>
#include <tuple>

namespace dcegm {

static const double kPi = 3.141592653589793;

void linrange(double lo, double hi, int n, double* out) {
    if (n <= 0) return;
    if (n == 1) {
        out[0] = lo;
        return;
    }
    const long double l = lo, h = hi, d = (long double)(n - 1);
    for (int i = 0; i < n; i++) out[i] = (double)((l * (long double)(n - 1 - i) + h * (long double)i) / d);
    out[0] = lo;
    out[n - 1] = hi;
}

// utils.jl:261-297: Newton on the Legendre recurrence, tolerance 3e-14, same starting guess
void gauss_legendre01(int n, std::vector<double>& x, std::vector<double>& w) {
    x.assign(n, 0.0);
    w.assign(n, 0.0);
    const double lo = 0.0, hi = 1.0;
    const double xm = (hi + lo) / 2, xl = (hi - lo) / 2;
    const int half = (int)std::floor((n + 1) / 2.0);
    double zprev = 1.e99, dp = 0;
    for (int i = 1; i <= half; i++) {
        double z = std::cos(kPi * (i - 0.25) / (n + 0.5));
        while (std::fabs(z - zprev) > 3.e-14) {
            double pa = 1, pb = 0;
            for (int j = 1; j <= n; j++) {
                double pc = pb;
                pb = pa;
                pa = ((2 * j - 1) * z * pb - (j - 1) * pc) / j;
            }
            dp = n * (z * pa - pb) / (z * z - 1);
            zprev = z;
            z = zprev - pa / dp;
        }
        x[i - 1] = xm - xl * z;
        x[n - i] = xm + xl * z;
        w[i - 1] = 2 * xl / ((1 - z * z) * dp * dp);
        w[n - i] = w[i - 1];
    }
}

// Wichura's AS241 (PPND16) + one Halley step against erfc
double normal_quantile(double p) {
    if (!(p > 0.0)) return -std::numeric_limits<double>::infinity();
    if (!(p < 1.0)) return std::numeric_limits<double>::infinity();
    static const double a[8] = {3.3871328727963666080e0, 1.3314166789178437745e2, 1.9715909503065514427e3,
                                1.3731693765509461125e4, 4.5921953931549871457e4, 6.7265770927008700853e4,
                                3.3430575583588128105e4, 2.5090809287301226727e3};
    static const double b[8] = {1.0, 4.2313330701600911252e1, 6.8718700749205790830e2, 5.3941960214247511077e3,
                                2.1213794301586595867e4, 3.9307895800092710610e4, 2.8729085735721942674e4,
                                5.2264952788528545610e3};
    static const double c[8] = {1.42343711074968357734e0, 4.63033784615654529590e0, 5.76949722146069140550e0,
                                3.64784832476320460504e0, 1.27045825245236838258e0, 2.41780725177450611770e-1,
                                2.27238449892691845833e-2, 7.74545014278341407640e-4};
    static const double d[8] = {1.0, 2.05319162663775882187e0, 1.67638483018380384940e0, 6.89767334985100004550e-1,
                                1.48103976427480074590e-1, 1.51986665636164571966e-2, 5.47593808499534494600e-4,
                                1.05075007164441684324e-9};
    static const double e[8] = {6.65790464350110377720e0, 5.46378491116411436990e0, 1.78482653991729133580e0,
                                2.96560571828504891230e-1, 2.65321895265761230930e-2, 1.24266094738807843860e-3,
                                2.71155556874348757815e-5, 2.01033439929228813265e-7};
    static const double f[8] = {1.0, 5.99832206555887937690e-1, 1.36929880922735805310e-1, 1.48753612908506148525e-2,
                                7.86869131145613259100e-4, 1.84631831751005468180e-5, 1.42151175831644588870e-7,
                                2.04426310338993978564e-15};
    auto horner = [](const double* k, double r) {
        double s = k[7];
        for (int i = 6; i >= 0; i--) s = s * r + k[i];
        return s;
    };
    const double q = p - 0.5;
    double val;
    if (std::fabs(q) <= 0.425) {
        const double r = 0.180625 - q * q;
        val = q * horner(a, r) / horner(b, r);
    } else {
        double r = std::sqrt(-std::log(q < 0 ? p : 1.0 - p));
        if (r <= 5.0) {
            r -= 1.6;
            val = horner(c, r) / horner(d, r);
        } else {
            r -= 5.0;
            val = horner(e, r) / horner(f, r);
        }
        if (q < 0) val = -val;
    }
    const double err = 0.5 * std::erfc(-val / std::sqrt(2.0)) - p;
    const double t = err * std::sqrt(2.0 * kPi) * std::exp(val * val / 2.0);
    return val - t / (1.0 + val * t / 2.0);
}

// Newton on the orthonormal Hermite recurrence in long double; nodes ascending.
void gauss_hermite(int n, std::vector<double>& x, std::vector<double>& w) {
    x.assign(n, 0.0);
    w.assign(n, 0.0);
    const long double pim4 = 0.7511255444649425L;
    const int half = (n + 1) / 2;
    std::vector<long double> r(half);
    long double z = 0, dp = 1;
    for (int i = 0; i < half; i++) {
        switch (i) {
            case 0: z = sqrtl((long double)(2 * n + 1)) - 1.85575L * powl((long double)(2 * n + 1), -0.16667L); break;
            case 1: z -= 1.14L * powl((long double)n, 0.426L) / z; break;
            case 2: z = 1.86L * z - 0.86L * r[0]; break;
            case 3: z = 1.91L * z - 0.91L * r[1]; break;
            default: z = 2.0L * z - r[i - 2];
        }
        for (int iter = 0; iter < 200; iter++) {
            long double pa = pim4, pb = 0.0L;
            for (int j = 0; j < n; j++) {
                long double pc = pb;
                pb = pa;
                pa = z * sqrtl(2.0L / (j + 1)) * pb - sqrtl((long double)j / (j + 1)) * pc;
            }
            dp = sqrtl((long double)(2 * n)) * pb;
            long double zo = z;
            z = zo - pa / dp;
            if (fabsl(z - zo) <= 1e-19L * (1.0L + fabsl(z))) break;
        }
        r[i] = z;
        x[n - 1 - i] = (double)z;
        x[i] = (double)(-z);
        w[i] = w[n - 1 - i] = (double)(2.0L / (dp * dp));
    }
    if (n % 2 == 1) x[n / 2] = 0.0;
}

void rouwenhorst(double rho, double mu, double sigma, int n, std::vector<double>& z, std::vector<double>& P) {
    const double q = (rho + 1) / 2;
    const double spread = std::sqrt((n - 1) / (1 - rho * rho)) * sigma;
    std::vector<double> cur = {q, 1 - q, 1 - q, q};
    int sz = 2;
    for (int i = 2; i <= n - 1; i++) {
        const int ns = i + 1;
        std::vector<double> nxt((size_t)ns * ns, 0.0);
        auto old = [&](int r, int c) { return (r < 0 || c < 0 || r >= sz || c >= sz) ? 0.0 : cur[(size_t)r * sz + c]; };
        for (int r = 0; r < ns; r++)
            for (int c = 0; c < ns; c++) {
                // q*[P 0;0 0] .+ (1-q)*[0 P;0 0] .+ (1-q)*[0 0;P 0] .+ q*[0 0;0 P], left to right
                double v = q * old(r, c) + (1 - q) * old(r, c - 1);
                v = v + (1 - q) * old(r - 1, c);
                v = v + q * old(r - 1, c - 1);
                nxt[(size_t)r * ns + c] = v;
            }
        for (int r = 1; r <= i - 1; r++)
            for (int c = 0; c < ns; c++) nxt[(size_t)r * ns + c] /= 2;
        cur.swap(nxt);
        sz = ns;
    }
    P = cur;
    z.resize(n);
    linrange(mu / (1 - rho) - spread, mu / (1 - rho) + spread, n, z.data());
}

void scale_grid(double lb, double ub, int n, int logorder, double* out) {
    if (lb >= ub) lb = ub - std::fabs(ub) * 0.05;
    const double off = lb < 0 ? 1 - lb : 1;
    auto fwd = [&](double v) { return std::log(v + off); };
    auto bwd = [&](double v) { return std::exp(v) - off; };
    double a = lb, b = ub;
    for (int k = 0; k < logorder; k++) {
        a = fwd(a);
        b = fwd(b);
    }
    linrange(a, b, n, out);
    for (int i = 0; i < n; i++) {
        double v = out[i];
        for (int k = 0; k < logorder; k++) v = bwd(v);
        out[i] = v;
    }
}

double income(int it, const dcegm_param& p, double shock) {
    const int age = it + 19;
    if (it > p.retage) return p.pension;
    return std::exp(p.inc0 + p.inc1 * age - p.inc2 * (double)(age * age) + shock);
}

static DevModel make_dev(const dcegm_param& p) {
    DevModel d{};
    d.gamma = p.gamma;
    d.beta = p.beta;
    d.R = p.R;
    d.lambda = p.lambda;
    d.cfloor = p.cfloor;
    d.alpha = p.alpha;
    d.delta = p.delta;
    d.nu = p.nu;
    d.bbar = p.bbar;
    d.a_low = p.a_low;
    d.a_lowT = p.a_lowT;
    d.a_high = p.a_high;
    d.pension = p.pension;
    d.neg_gamma = (-1.0) * p.gamma;
    d.oneminusgamma = 1.0 - p.gamma;
    d.oneover_oneminusgamma = 1.0 / d.oneminusgamma;
    d.neg_oneover_gamma = (-1.0) / p.gamma;
    d.betaR = p.beta * p.R;
    d.betaRnu = p.beta * p.R * p.nu;
    d.na = p.na;
    d.ny = p.ny;
    d.nT = p.nT;
    d.quad_group = p.quad_group > 0 ? p.quad_group : 1;
    d.gamma_is_one = p.gamma == 1.0;
    return d;
}

bool build_tables(int model, const dcegm_param* params, int nmodels, HostTables& out, const char** why) {
    static const char* kWhy[] = {"nmodels must be positive", "all models of one call must share na, ny and nT",
                                 "na >= 2, ny >= 1, nT >= 2, nD == 2 required", "quad_group > 1 (lane-group quadrature sums) is not implemented by the kernels: pass 0 or 1",
                                 "unknown model kind"};
    if (nmodels <= 0 || !params) {
        *why = kWhy[0];
        return false;
    }
    const int na = params[0].na, ny = params[0].ny, nT = params[0].nT;
    for (int m = 0; m < nmodels; m++) {
        const dcegm_param& p = params[m];
        if (p.na != na || p.ny != ny || p.nT != nT) {
            *why = kWhy[1];
            return false;
        }
        if (p.na < 2 || p.ny < 1 || p.nT < 2 || p.nD != 2) {
            *why = kWhy[2];
            return false;
        }
        if (p.quad_group > 1) {  // the kernels sum the nodes in ascending serial order
            *why = kWhy[3];
            return false;
        }
    }
    const bool minimal = model == DCEGM_MODEL_MINIMAL || model == DCEGM_MODEL_MINIMAL_BORROW || model == DCEGM_MODEL_MINIMAL_BEQUEST;
    if (model != DCEGM_MODEL_F && model != DCEGM_MODEL_G && !minimal) {
        *why = kWhy[4];
        return false;
    }

    out.models.resize(nmodels);
    // savings grids, shared between models with the same bounds
    std::map<std::tuple<double, double>, int> grid_of;
    out.grids.clear();
    out.ngrids = 0;
    for (int m = 0; m < nmodels; m++) {
        const dcegm_param& p = params[m];
        out.models[m] = make_dev(p);
        // minimal_EGM.jl:7 (from 0), :53 (from a_low), :97 (scaleGrid from 0)
        const double glo = (model == DCEGM_MODEL_MINIMAL || model == DCEGM_MODEL_MINIMAL_BEQUEST) ? 0.0 : p.a_low;
        auto key = std::make_tuple(glo, p.a_high);
        auto it = grid_of.find(key);
        if (it == grid_of.end()) {
            out.grids.resize((size_t)(out.ngrids + 1) * na);
            double* g = out.grids.data() + (size_t)out.ngrids * na;
            if (model == DCEGM_MODEL_G)
                scale_grid(p.a_low, p.a_high, na, 2, g);  // param.jl:288
            else if (model == DCEGM_MODEL_MINIMAL_BEQUEST)
                scale_grid(0.0, p.a_high, na, 1, g);      // minimal_EGM.jl:97
            else
                linrange(glo, p.a_high, na, g);  // param.jl:205
            it = grid_of.emplace(key, out.ngrids++).first;
        }
        out.models[m].grid = it->second;
    }

    // quadrature rules depend on ny only
    std::vector<double> glx, glw, ghx, ghw, zq(ny);
    if (model == DCEGM_MODEL_F) {
        gauss_legendre01(ny, glx, glw);                                  // param.jl:194
        for (int i = 0; i < ny; i++) zq[i] = normal_quantile(glx[i]);   // param.jl:195-196
    } else {
        gauss_hermite(ny, ghx, ghw);  // param.jl:279, minimal_EGM.jl:4
    }
    const double inv_sqrt_pi = std::pow(kPi, -0.5);

    const size_t off_per = minimal ? (size_t)ny : (size_t)(nT - 1) * 2 * ny;
    const size_t wgt_per = model == DCEGM_MODEL_G ? (size_t)ny * ny : (size_t)ny;
    // every element below is written by one(m): no zero fill (7.9 MB of offsets at 4096 C3 models)
    out.sim.resize(nmodels);
    out.yvec.resize((size_t)ny * nmodels);
    out.off.resize(off_per * nmodels);
    out.wgt.resize(wgt_per * nmodels);

    auto one = [&](int m) {
        const dcegm_param& p = params[m];
        double* off = out.off.data() + off_per * m;
        double* wgt = out.wgt.data() + wgt_per * m;
        double* yvec = out.yvec.data() + (size_t)ny * m;
        SimModel& sm = out.sim[m];
        sm = SimModel{};
        sm.inc0 = p.inc0;
        sm.inc1 = p.inc1;
        sm.inc2 = p.inc2;
        sm.sigma = p.sigma;
        sm.initw0 = p.initw0;
        sm.initw1 = p.initw1;
        sm.rseed = p.rseed;
        sm.retage = p.retage;
        sm.nsims = p.nsims;
        if (model == DCEGM_MODEL_F) {
            for (int i = 0; i < ny; i++) {
                yvec[i] = zq[i] * p.sigma;  // param.jl:197
                wgt[i] = glw[i];
            }
            for (int it = 1; it <= nT - 1; it++)  // reference period it, income of it+1
                for (int iy = 0; iy < ny; iy++) {
                    const double inc = income(it + 1, p, yvec[iy]);
                    off[((size_t)(it - 1) * 2 + 0) * ny + iy] = inc;                                        // id=1
                    off[((size_t)(it - 1) * 2 + 1) * ny + iy] = p.delta * inc + (1 - p.delta) * p.pension;  // id=2
                }
        } else if (model == DCEGM_MODEL_G) {
            if (p.rho == 0) {  // param.jl:278-281
                for (int i = 0; i < ny; i++) yvec[i] = std::sqrt(2.0) * p.sigma * ghx[i];
                for (int iy = 0; iy < ny; iy++)
                    for (int jy = 0; jy < ny; jy++) wgt[(size_t)iy * ny + jy] = ghw[jy] * inv_sqrt_pi;
            } else {  // param.jl:284
                std::vector<double> z, P;
                rouwenhorst(p.rho, 0.0, p.sigma, ny, z, P);
                std::copy(z.begin(), z.end(), yvec);
                std::copy(P.begin(), P.end(), wgt);
            }
            for (int it = 1; it <= nT - 1; it++)  // m1[it+1][iid][jy,:]  param.jl:301
                for (int jy = 0; jy < ny; jy++) {
                    off[((size_t)(it - 1) * 2 + 0) * ny + jy] = income(it + 1, p, yvec[jy]);
                    off[((size_t)(it - 1) * 2 + 1) * ny + jy] = p.pension;
                }
        } else {
            for (int i = 0; i < ny; i++) {
                yvec[i] = std::sqrt(2.0) * p.sigma * ghx[i];  // minimal_EGM.jl:5
                wgt[i] = ghw[i] * inv_sqrt_pi;                // minimal_EGM.jl:6
                off[i] = 0.0 + std::exp(yvec[i]);             // minimal_EGM.jl:21
            }
        }
    };
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<unsigned>(hw ? hw : 1, 16);
    if (nmodels < 256) nt = 1;
    if (nt <= 1) {
        for (int m = 0; m < nmodels; m++) one(m);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++)
            pool.emplace_back([&, t]() {
                for (int m = t; m < nmodels; m += nt) one(m);
            });
        for (auto& th : pool) th.join();
    }
    return true;
}

}  // namespace dcegm
