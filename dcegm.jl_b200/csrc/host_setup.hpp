// host_setup.hpp — host-side model constructors of the CUDA library.
//
// Replaces FModel(p) (src/param.jl:189-223), GModel(p) (param.jl:261-310) and the setup lines of
// minimal_EGM (src/minimal_EGM.jl:4-7): quadrature nodes/weights, savings grids and the
// next-period income offsets.  These are O(ny*nT + na) scalars per model; they are computed on
// the host in IEEE double (glibc libm) and uploaded, so that the kernels receive tables rather
// than re-deriving them with device transcendentals.
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/dcegm.h"

namespace dcegm {

// range(lo, stop=hi, length=n) of Julia (twice-precision StepRangeLen): correctly rounded
// lo + i*(hi-lo)/(n-1), evaluated in long double.
void linrange(double lo, double hi, int n, double* out);
// utils.jl:261-297
void gauss_legendre01(int n, std::vector<double>& x, std::vector<double>& w);
// quantile(Normal(0,1), p)
double normal_quantile(double p);
// FastGaussQuadrature.gausshermite
void gauss_hermite(int n, std::vector<double>& x, std::vector<double>& w);
// utils.jl:178-197; P row-major
void rouwenhorst(double rho, double mu, double sigma, int n, std::vector<double>& z, std::vector<double>& P);
// utils.jl:200-238
void scale_grid(double lb, double ub, int n, int logorder, double* out);
// utils.jl:243-250 (it is 1-based)
double income(int it, const dcegm_param& p, double shock);

// Device-side view of one model: scalars of Param + derived fields (param.jl:90-93).
struct DevModel {
    double gamma, beta, R, lambda, cfloor, alpha, delta, nu, bbar;
    double a_low, a_lowT, a_high, pension;
    double neg_gamma, oneminusgamma, oneover_oneminusgamma, neg_oneover_gamma;
    double betaR;        // beta*R
    double betaRnu;      // beta*R*nu  (bequest period, dc_algo.jl:50)
    int32_t na, ny, nT;
    int32_t grid;        // index into the savings-grid table
    int32_t quad_group;
    int32_t gamma_is_one;
    int32_t _pad[2];
};

// Scalars only the forward simulator needs (simulate.jl:61-203, utils.jl:243-250)
struct SimModel {
    double inc0, inc1, inc2, sigma, initw0, initw1;
    int64_t rseed;
    int32_t retage, nsims;
};

struct HostTables {
    std::vector<DevModel> models;
    std::vector<SimModel> sim;
    std::vector<double> yvec;    // G: [nmodels][ny] income support (param.jl:279-285); F/minimal: sigma-scaled nodes
    std::vector<double> grids;   // [ngrids][na]
    int ngrids = 0;
    // F: off [nmodels][nT-1][2 id][ny]   m1 = a*R + off          (param.jl:212-213)
    // G: off [nmodels][nT-1][2 iid][ny]  m1 = a*R + off          (param.jl:301), period it+1
    // minimal: off [nmodels][ny] = 0.0 + exp(yvec)               (minimal_EGM.jl:21)
    std::vector<double> off;
    // F/minimal: wgt [nmodels][ny];  G: wgt [nmodels][ny][ny] (row iy, col jy)
    std::vector<double> wgt;
};

// Builds all tables for a batch.  model = DCEGM_MODEL_*.  Returns false on invalid input.
bool build_tables(int model, const dcegm_param* params, int nmodels, HostTables& out, const char** why);

}  // namespace dcegm
