/* dc_rng.h — the shock source of the forward simulator: counter-based, stateless, reproducible.
 *
 * The reference seeds Julia's global RNG (Random.seed!(p.rseed), src/simulate.jl:72,137) and draws
 * whole matrices; its stream cannot be reproduced outside Julia.  Here every draw is a pure function
 * of (seed, agent, period, stream): Philox4x32-10 (Salmon et al., SC'11) keyed by the 64-bit seed,
 * counter = (agent, period, stream >> 1, 0), words 0-1 for even streams and 2-3 for odd ones.  Like the reference, the model index does not enter — two
 * parameterisations simulated with one seed see the same shocks (common random numbers).
 *
 * Integer arithmetic only for the bits; the uniform -> normal map is Wichura's AS241 (PPND16, relative
 * accuracy ~1e-16) written with IEEE operations, dc_log and sqrt, so the GPU kernel and the CPU oracle
 * produce identical shocks bit for bit.
 */
#ifndef DC_RNG_H
#define DC_RNG_H

#include <math.h>

#include "dc_math.h"

/* streams (simulate.jl:73-76, 138-141) */
#define DC_STREAM_W 0  /* wshocks  */
#define DC_STREAM_D 1  /* dshocks  */
#define DC_STREAM_R 2  /* rshocks  */
#define DC_STREAM_W0 3 /* w0shocks (period 0) */
#define DC_STREAM_Y0 4 /* initial income state (period 0, GModel) */

DC_HD uint32_t dc_mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

DC_HD void dc_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out) {
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = dc_mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = dc_mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0;
    out[1] = c1;
    out[2] = c2;
    out[3] = c3;
}

/* Uniforms on [0,1) with 53 random bits, like Julia's rand(Float64).  One Philox block (four words) serves two
 * streams: block = stream >> 1, the even stream takes words 0-1 and the odd one words 2-3 — so the simulator gets an
 * agent's income shock (W) and choice shock (D) of a period from one call. */
DC_HD double dc_bits53(uint32_t a, uint32_t b) {
    const uint64_t bits = ((uint64_t)(a >> 5) << 26) | (uint64_t)(b >> 6);
    return (double)(int64_t)bits * 1.1102230246251565e-16; /* 2^-53, exact */
}
DC_HD void dc_sim_uniform_pair(int64_t seed, uint32_t agent, uint32_t period, uint32_t block, double* even, double* odd) {
    uint32_t o[4];
    dc_philox4x32_10(agent, period, block, 0u, (uint32_t)((uint64_t)seed & 0xffffffffu), (uint32_t)((uint64_t)seed >> 32), o);
    *even = dc_bits53(o[0], o[1]);
    *odd = dc_bits53(o[2], o[3]);
}
DC_HD double dc_sim_uniform(int64_t seed, uint32_t agent, uint32_t period, uint32_t stream) {
    double e, o;
    dc_sim_uniform_pair(seed, agent, period, stream >> 1, &e, &o);
    return (stream & 1u) ? o : e;
}

DC_HD double dc_horner8(const double* k, double r) {
    double s = k[7];
    for (int i = 6; i >= 0; i--) s = s * r + k[i];
    return s;
}

/* quantile(Normal(0,1), p): AS241 PPND16.  p <= 0 -> -Inf, p >= 1 -> +Inf (Distributions.jl does the same) */
DC_HD double dc_norm_quantile(double p) {
    const double a[8] = {3.3871328727963666080e0, 1.3314166789178437745e2, 1.9715909503065514427e3, 1.3731693765509461125e4,
                         4.5921953931549871457e4, 6.7265770927008700853e4, 3.3430575583588128105e4, 2.5090809287301226727e3};
    const double b[8] = {1.0, 4.2313330701600911252e1, 6.8718700749205790830e2, 5.3941960214247511077e3,
                         2.1213794301586595867e4, 3.9307895800092710610e4, 2.8729085735721942674e4, 5.2264952788528545610e3};
    const double c[8] = {1.42343711074968357734e0, 4.63033784615654529590e0, 5.76949722146069140550e0, 3.64784832476320460504e0,
                         1.27045825245236838258e0, 2.41780725177450611770e-1, 2.27238449892691845833e-2, 7.74545014278341407640e-4};
    const double d[8] = {1.0, 2.05319162663775882187e0, 1.67638483018380384940e0, 6.89767334985100004550e-1,
                         1.48103976427480074590e-1, 1.51986665636164571966e-2, 5.47593808499534494600e-4, 1.05075007164441684324e-9};
    const double e[8] = {6.65790464350110377720e0, 5.46378491116411436990e0, 1.78482653991729133580e0, 2.96560571828504891230e-1,
                         2.65321895265761230930e-2, 1.24266094738807843860e-3, 2.71155556874348757815e-5, 2.01033439929228813265e-7};
    const double f[8] = {1.0, 5.99832206555887937690e-1, 1.36929880922735805310e-1, 1.48753612908506148525e-2,
                         7.86869131145613259100e-4, 1.84631831751005468180e-5, 1.42151175831644588870e-7, 2.04426310338993978564e-15};
    if (!(p > 0.0)) return -DC_INF;
    if (!(p < 1.0)) return DC_INF;
    const double q = p - 0.5;
    if (fabs(q) <= 0.425) {
        const double r = 0.180625 - q * q;
        return q * dc_horner8(a, r) / dc_horner8(b, r);
    }
    double r = sqrt(-dc_log(q < 0 ? p : 1.0 - p));
    double val;
    if (r <= 5.0) {
        r -= 1.6;
        val = dc_horner8(c, r) / dc_horner8(d, r);
    } else {
        r -= 5.0;
        val = dc_horner8(e, r) / dc_horner8(f, r);
    }
    return q < 0 ? -val : val;
}

#endif /* DC_RNG_H */
