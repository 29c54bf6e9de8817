// solver_sim.cu — forward panel simulation of the resident solution.
//
// Replaces sim(m::FModel,p) (src/simulate.jl:61-130) and sim(m::GModel,p) (simulate.jl:133-203).
// One CTA per model.  Agents are processed in tiles of kSimTile (4 per thread, state in registers);
// for every tile the CTA walks the periods forward, staging that period's policy/value polylines
// (all income states) from HBM/L2 into shared memory as SoA, so nothing per-agent is ever written to
// memory unless the caller asked for the panel.  Shocks are counter-based (dc_rng.h) or given.
// Per-period sums are reduced tile by tile in a fixed order (the CPU oracle reproduces it).
#include <math_constants.h>


#include "dc_rng.h"
#include "device_common.cuh"

namespace dcegm {

// unrolling of the per-agent body inside one period: each copy is ~1500 instructions and four of them do not fit the
// instruction cache (no_instruction was 19 % of the stall samples).  Rolled, the agents' state arrays are indexed
// dynamically (a few local-memory accesses per agent-period, L1 hits).  Measured, G / F models:
// unroll 4: 13.9 / 14.5, unroll 2: 14.7 / 15.7, rolled: 16.1 / 18.3 G agent-periods/s.
#ifndef DC_SIM_UNROLL
#define DC_SIM_UNROLL 1
#endif
constexpr int kSimUnroll = DC_SIM_UNROLL;
constexpr int kSimAgents = 4;                    // agents per thread and tile
constexpr int kSimTile = kThreads * kSimAgents;  // agents per tile: the unit of the fixed summation order
// tiles that share one staging of a period's lines (their sums are still formed and added tile by tile)
// (measured, G agent-periods/s: 16.5 / 17.2 / 18.5 / 19.5 for 1 / 2 / 4 / 8 tiles with five income states staged per
// period; with the single state of the retirement model staging is cheap and one tile is fastest: 19.0 vs 16.7)
constexpr int kSimSubMax = 8;
__host__ __device__ constexpr int sim_sub(int model) { return model == DCEGM_MODEL_G ? 8 : 1; }


// global_lines: the variant that reads the polylines in place (HBM / L2) instead of staging them per period — for
// lines too long for shared memory
size_t sim_smem_bytes(const dcegm_layout& l, bool global_lines) {
    const size_t lines = global_lines ? 0 : (size_t)l.nstate * 8 * l.cap * sizeof(double);
    const size_t tabs = ((size_t)l.ny * l.ny + (size_t)l.nT * l.ny + (size_t)l.nT * DCEGM_SIM_NMOM +
                         (size_t)kSimSubMax * kWarps * DCEGM_SIM_NMOM +
                         (size_t)l.nstate * 2) * sizeof(double);
    const size_t ints = ((size_t)l.nT + 1 + (size_t)l.nstate * 4 + 4) * sizeof(int);
    return lines + tabs + ints + 64;
}

// vfun (dc_algo.jl:220-251) on the value line of period `it` (0-based)
template <int Top>
__device__ __forceinline__ double sim_vfun(bool worker_line, bool terminal, double c1, double m1, const double* vx,
                                           const double* vy, int n, double vb, const DevModel& p) {
    if (terminal && p.nu > 0) return p.nu * u_plain(c1 + p.bbar, p) - 0.0;  // bequest(c1,false,p)
    const bool mask = terminal || (m1 < vx[1]);
    if (mask) return u_work(c1, worker_line, p) + p.beta * vb;
    return interp_y<Top>(vx, vy, n, m1);
}

// The same interpolation on a stored line (interleaved (x, y) points in HBM): the operations of interp_y / interp_at
// (device_common.cuh) in the same order, so both variants of the simulator give the same bits.
template <int Top>
__device__ __forceinline__ double interp_y_aos(const double2* __restrict__ L, int n, double q) {
    int lo = 0;
#pragma unroll
    for (int step = Top; step >= 1; step >>= 1) {
        const int idx = lo + step - 1;
        const bool ok = idx < n && L[idx < n ? idx : 0].x <= q;
        lo = ok ? lo + step : lo;
    }
    int i = lo - 1;
    i = i < 0 ? 0 : i;
    i = i > n - 2 ? n - 2 : i;
    const double2 p0 = L[0], p1 = L[1], pa = L[n - 2], pb = L[n - 1];
    if (q < p0.x) {
        double d = p1.x - p0.x;
        return p0.y + ((p1.y - p0.y) / d) * (q - p0.x);
    }
    if (q > pb.x) {
        double d = pb.x - pa.x;
        return pb.y + ((pb.y - pa.y) / d) * (q - pb.x);
    }
    const double2 a = L[i], b = L[i + 1];
    double f = (q - a.x) / (b.x - a.x);
    return (1.0 - f) * a.y + f * b.y;
}

template <int Top>
__device__ __forceinline__ double sim_vfun_aos(bool worker_line, bool terminal, double c1, double m1, const double2* L, int n,
                                               double vb, const DevModel& p) {
    if (terminal && p.nu > 0) return p.nu * u_plain(c1 + p.bbar, p) - 0.0;
    const bool mask = terminal || (m1 < L[1].x);
    if (mask) return u_work(c1, worker_line, p) + p.beta * vb;
    return interp_y_aos<Top>(L, n, m1);
}

// Top = largest step of the fixed-trip knot search: lines hold fewer than 2*Top points (chosen at launch from cap).
// GL: read the lines in place (global memory) instead of staging each period's lines into shared memory.
template <int MODEL, int Top, bool GL = false>
__global__ void __launch_bounds__(kThreads, 3) sim_kernel(DeviceBatch B, SimBatch Sb) {
    const int m = B.first + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    constexpr int kSimSub = sim_sub(MODEL);
    constexpr int kSimState = kSimSub * kSimAgents;  // agents per thread between two stagings
    __shared__ DevModel s_p;
    __shared__ SimModel s_sp;
    __shared__ int s_status;
    if (tid == 0) {
        s_p = B.models[m];
        s_sp = Sb.sim[m];
        s_status = 0;
    }
    __syncthreads();
    const DevModel& p = s_p;
    const SimModel& sp = s_sp;
    const dcegm_layout& L = B.lay;
    const int nT = L.nT, ny = L.ny, nstate = L.nstate, cap = L.cap;
    const int nsims = Sb.given ? Sb.nsims_given : sp.nsims;

    // shared-memory map
    double* ln = reinterpret_cast<double*>(dc_smem);                 // [nstate][id][kind][x|y][cap]
    double* gg = ln + (GL ? (size_t)0 : (size_t)nstate * 8 * cap);   // [ny][ny] cumulative transition rows
    double* inctab = gg + ny * ny;                                   // [nT][ny] income(it+1, yvec)
    double* mom = inctab + nT * ny;                                  // [nT][NMOM]
    double* red = mom + nT * DCEGM_SIM_NMOM;                         // [kSimSub][kWarps][NMOM] warp sums of one period
    double* vb = red + kSimSubMax * kWarps * DCEGM_SIM_NMOM;         // [nstate][2]
    int* hist = reinterpret_cast<int*>(vb + nstate * 2);             // [nT+1]
    int* lens = hist + nT + 1;                                       // [nstate][id][kind]
    auto line = [&](int st, int id, int kind, int xy) { return ln + ((((size_t)st * 2 + id) * 2 + kind) * 2 + xy) * cap; };

    for (int k = tid; k < nT + 1; k += kThreads) hist[k] = 0;
    for (int k = tid; k < nT * DCEGM_SIM_NMOM; k += kThreads) mom[k] = 0.0;
    if (MODEL == DCEGM_MODEL_G) {
        // GG = cumsum(m.ywgt, dims=2) (simulate.jl:145), sequential like Julia's cumsum
        const double* w = B.wgt + (size_t)m * ny * ny;
        for (int r = tid; r < ny; r += kThreads) {
            double acc = 0.0;
            for (int j = 0; j < ny; j++) {
                acc = j == 0 ? w[r * ny + j] : acc + w[r * ny + j];
                gg[r * ny + j] = acc;
            }
        }
        // income(it, p, yvec[iy]) (utils.jl:243-250), it 1-based
        const double* yv = Sb.yvec + (size_t)m * ny;
        for (int k = tid; k < nT * ny; k += kThreads) {
            const int it1 = k / ny + 1, iy = k % ny;
            const int age = it1 + 19;
            inctab[k] = it1 > sp.retage ? p.pension
                                        : dc_exp(sp.inc0 + sp.inc1 * (double)age - sp.inc2 * (double)(age * age) + yv[iy]);
        }
    }
    const double a_lo = p.a_low, a_hi = p.a_high;
    const size_t npan = (size_t)nsims * nT;
    double* pd = Sb.pan_d ? Sb.pan_d + (size_t)m * 6 * npan : nullptr;
    int32_t* pi = Sb.pan_i ? Sb.pan_i + (size_t)m * 2 * npan : nullptr;

    // warp sums of period `itf` (in red) -> running per-period sums, tile by tile in ascending order, warps ascending
    auto fold = [&](int itf, int group0) {
        if (tid < DCEGM_SIM_NMOM) {
            for (int lt = 0; lt < kSimSub; lt++) {
                const int t0 = group0 + lt * kSimTile;
                if (t0 >= nsims) break;
                double t = red[(lt * kWarps) * DCEGM_SIM_NMOM + tid];
                for (int w = 1; w < kWarps; w++) t += red[(lt * kWarps + w) * DCEGM_SIM_NMOM + tid];
                mom[itf * DCEGM_SIM_NMOM + tid] = t0 == 0 ? t : mom[itf * DCEGM_SIM_NMOM + tid] + t;
            }
        }
    };
    for (int group0 = 0; group0 < nsims; group0 += kSimTile * kSimSub) {
        // the agents' state between periods lives in local memory (the per-agent body is rolled): wealth, and one
        // word holding the income state (bits 0-7), "works this period" (bit 8) and the retirement period (bits 16-31)
        double w1[kSimState];
        unsigned stw[kSimState];
#pragma unroll 1
        for (int a = 0; a < kSimState; a++) {
            w1[a] = 0.0;
            stw[a] = 1u << 8;
        }
        for (int it = 0; it < nT; it++) {
            __syncthreads();
            if (it > 0) fold(it - 1, group0);  // eight threads; the others go on staging
            // ---- stage the lines of period `it` (AoS in HBM -> SoA in shared memory) ----
            for (int s = tid; s < nstate * 4; s += kThreads) {
                const size_t slot = ((size_t)m * nT + it) * nstate * 4 + s;
                const int n = B.lens[slot];
                lens[s] = n < cap ? n : cap;
            }
            for (int s = tid; s < nstate * 2; s += kThreads) vb[s] = B.vbound[((size_t)m * nT + it) * nstate * 2 + s];
            if (!GL) {   // all (slot, point) pairs of the period in one flat loop: independent loads, no per-slot chain
                const size_t slot0 = ((size_t)m * nT + it) * nstate * 4;
                const double2* src0 = B.lines + slot0 * (size_t)cap;
                const int32_t* len0 = B.lens + slot0;
                const int total = nstate * 4 * cap;
#pragma unroll 4
                for (int idx = tid; idx < total; idx += kThreads) {
                    const int s = idx / cap, k = idx - s * cap;
                    if (k < len0[s]) {
                        const double2 v = src0[idx];
                        double* dx = ln + ((size_t)s * 2) * cap;
                        dx[k] = v.x;
                        dx[cap + k] = v.y;
                    }
                }
            }
            __syncthreads();
            int flags = 0;
#pragma unroll 1
            for (int lt = 0; lt < kSimSub; lt++) {
            const int tile0 = group0 + lt * kSimTile;
            if (tile0 >= nsims) break;  // block-uniform
            double sum[DCEGM_SIM_NMOM];
#pragma unroll
            for (int k = 0; k < DCEGM_SIM_NMOM; k++) sum[k] = 0.0;
            // the next agent's state is fetched from local memory one iteration ahead (ncu: 8 % of the samples waited
            // for these two loads at their first use; G +4 %.  Not for the retirement model, whose body is at the
            // register limit already: -1 %)
            constexpr bool kAhead = MODEL == DCEGM_MODEL_G;
            double w1_next = kAhead ? w1[lt * kSimAgents] : 0.0;
            unsigned sw_next = kAhead ? stw[lt * kSimAgents] : 0u;
#pragma unroll kSimUnroll
            for (int a0 = 0; a0 < kSimAgents; a0++) {
                const int a = lt * kSimAgents + a0;  // slot of the agent's state
                const int i = tile0 + a0 * kThreads + tid;
                const double w1a = kAhead ? w1_next : w1[a];
                const unsigned sw = kAhead ? sw_next : stw[a];
                if (kAhead && a0 + 1 < kSimAgents) {
                    w1_next = w1[a + 1];
                    sw_next = stw[a + 1];
                }
                if (i >= nsims) continue;
                const size_t at = (size_t)i + (size_t)nsims * it;
                // ---- state at the start of the period (simulate.jl:82-101, 160-178) ----
                double w0, inc;
                const bool wk = (sw >> 8) & 1u;
                int ysa = (int)(sw & 0xffu), reta = (int)(sw >> 16);
                const int worker = wk ? 1 : 2;
                // income shock and choice shock of this agent-period: one Philox block (dc_rng.h); nobody needs
                // either for a retiree of the retirement model
                double u_w = 0.0, u_d = 1.0;
                static_assert(DC_STREAM_W == 0 && DC_STREAM_D == 1, "W and D share block 0");
                if (!Sb.given && (wk || (MODEL == DCEGM_MODEL_G && it > 0)))
                    dc_sim_uniform_pair(sp.rseed, (uint32_t)i, (uint32_t)it, 0u, &u_w, &u_d);
                if (MODEL == DCEGM_MODEL_G) {
                    if (it == 0) {
                        int y0;
                        if (Sb.given)
                            y0 = Sb.y0[i] - 1;
                        else
                            y0 = (int)(dc_sim_uniform(sp.rseed, (uint32_t)i, 0u, DC_STREAM_Y0) * (double)ny);
                        y0 = y0 < 0 ? 0 : (y0 > ny - 1 ? ny - 1 : y0);
                        ysa = y0;
                    } else {
                        // searchsortedfirst(GG[ystate[i,it-1],:], wshocks[i,it]) (simulate.jl:148)
                        const double uw = Sb.given ? Sb.wsh[at] : u_w;
                        const double* row = gg + ysa * ny;
                        int j = 0;
                        while (j < ny && !(row[j] >= uw)) j++;
                        if (j >= ny) {
                            flags |= DCEGM_ST_YSTATE_RANGE;
                            j = ny - 1;
                        }
                        ysa = j;
                    }
                }
                if (it == 0) {
                    const double u0 = Sb.given ? Sb.w0sh[i] : dc_sim_uniform(sp.rseed, (uint32_t)i, 0u, DC_STREAM_W0);
                    w0 = sp.initw0 + u0 * (sp.initw1 - sp.initw0);
                    inc = MODEL == DCEGM_MODEL_G ? inctab[ysa] : __longlong_as_double(0x7ff8000000000000LL);  // stays NaN, simulate.jl:18
                } else {
                    if (!wk) {
                        inc = p.pension;
                    } else if (MODEL == DCEGM_MODEL_G) {
                        inc = inctab[it * ny + ysa];
                    } else {
                        const double uw = Sb.given ? Sb.wsh[at] : u_w;
                        const double shock = dc_norm_quantile(uw) * sp.sigma;  // simulate.jl:73
                        const int it1 = it + 1, age = it1 + 19;
                        inc = it1 > sp.retage ? p.pension
                                              : dc_exp(sp.inc0 + sp.inc1 * (double)age - sp.inc2 * (double)(age * age) + shock);
                    }
                    w0 = w1a * p.R + inc;  // simulate.jl:103
                }
                // ---- consumption, utility, end-of-period wealth (simulate.jl:106-110) ----
                const int st = MODEL == DCEGM_MODEL_G ? ysa : 0;
                const int id = wk ? 0 : 1;
                // lines of this period in HBM (GL variant): slot (st, id, kind)
                const double2* gl0 = B.lines + (((size_t)m * nT + it) * nstate * 4) * (size_t)cap;
                auto gline = [&](int st_, int id_, int kind_) { return gl0 + (size_t)((st_ * 2 + id_) * 2 + kind_) * cap; };
                const double cons = GL ? interp_y_aos<Top>(gline(st, id, 0), lens[(st * 2 + id) * 2 + 0], w0)
                                       : interp_y<Top>(line(st, id, 0, 0), line(st, id, 0, 1), lens[(st * 2 + id) * 2 + 0], w0);
                const double util = u_work(cons, wk, p);
                double w1n = w0 - cons;
                if (MODEL == DCEGM_MODEL_F) w1n = w1n > a_hi ? a_hi : (w1n < a_lo ? a_lo : w1n);  // clamp!, simulate.jl:109
                const bool negc = cons < 0.0;
                if (negc && MODEL == DCEGM_MODEL_F) flags |= DCEGM_ST_NEG_CONS;
                // ---- probability to remain worker (simulate.jl:119-121) ----
                const bool term = it == nT - 1;
                double pw = 0.0;  // retirees: simulate.jl:121 sets 0 whatever the values are
                if (wk) {
                    double v1, v2;
                    if (GL) {
                        v1 = sim_vfun_aos<Top>(true, term, cons, w0, gline(st, 0, 1), lens[(st * 2 + 0) * 2 + 1], vb[st * 2 + 0], p);
                        v2 = sim_vfun_aos<Top>(false, term, cons, w0, gline(st, 1, 1), lens[(st * 2 + 1) * 2 + 1], vb[st * 2 + 1], p);
                    } else {
                        v1 = sim_vfun<Top>(true, term, cons, w0, line(st, 0, 1, 0), line(st, 0, 1, 1),
                                           lens[(st * 2 + 0) * 2 + 1], vb[st * 2 + 0], p);
                        v2 = sim_vfun<Top>(false, term, cons, w0, line(st, 1, 1, 0), line(st, 1, 1, 1),
                                           lens[(st * 2 + 1) * 2 + 1], vb[st * 2 + 1], p);
                    }
                    // the larger value's exponential is exp(0) = 1 exactly: one dc_exp instead of two
                    const bool w_hi = v1 > v2;
                    const double mx = w_hi ? v1 : v2;
                    const double e = dc_exp(((w_hi ? v2 : v1) - mx) / p.lambda);
                    const double e1 = w_hi ? 1.0 : e, e2 = w_hi ? e : 1.0;
                    pw = e1 / (e1 + e2);
                }
                // ---- next period's state (simulate.jl:124-127, 201) ----
                // a retiree's draw cannot matter (pw = 0 is never above a uniform in [0,1)): no Philox call for it
                const double ud = Sb.given ? Sb.dsh[at] : (wk ? u_d : 1.0);
                bool next = pw > ud;
                if (MODEL == DCEGM_MODEL_F && !wk) {
                    const double ur = Sb.given ? (Sb.rsh ? Sb.rsh[at] : 1.0) : dc_sim_uniform(sp.rseed, (uint32_t)i, (uint32_t)it, DC_STREAM_R);
                    next = p.delta > ur;
                }
                // retirement period (simulate.jl:95 sets it at the start of the first period not worked, 1-based)
                if (wk && !next && it + 1 < nT) reta = it + 2;
                stw[a] = (unsigned)ysa | (next ? 1u << 8 : 0u) | ((unsigned)reta << 16);
                w1[a] = w1n;
                sum[0] += w0;
                sum[1] += w1n;
                sum[2] += cons;
                sum[3] += util;
                sum[4] += inc;
                sum[5] += pw;
                sum[6] += wk ? 1.0 : 0.0;
                sum[7] += negc ? 1.0 : 0.0;
                if (pd) {
                    pd[0 * npan + at] = w0;
                    pd[1 * npan + at] = w1n;
                    pd[2 * npan + at] = cons;
                    pd[3 * npan + at] = util;
                    pd[4 * npan + at] = inc;
                    pd[5 * npan + at] = pw;
                    pi[0 * npan + at] = MODEL == DCEGM_MODEL_G ? ysa + 1 : 0;
                    pi[1 * npan + at] = worker;
                }
            }
            // ---- tile sums: shuffle tree per warp; the warps are added up (ascending) by fold() ----
#pragma unroll
            for (int k = 0; k < DCEGM_SIM_NMOM; k++) {
                double v = sum[k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) red[(lt * kWarps + wid) * DCEGM_SIM_NMOM + k] = v;
            }
            }  // lt
            if (flags) atomicOr(&s_status, flags);
        }
        __syncthreads();
        fold(nT - 1, group0);
#pragma unroll 1
        for (int a = 0; a < kSimState; a++) {
            const int i = group0 + (a / kSimAgents) * kSimTile + (a % kSimAgents) * kThreads + tid;
            if (i < nsims) {
                const int reta = (int)(stw[a] >> 16);
                atomicAdd(&hist[reta], 1);
                if (Sb.pan_ret) Sb.pan_ret[(size_t)m * nsims + i] = reta;
            }
        }
    }
    __syncthreads();
    const double dn = (double)nsims;
    for (int k = tid; k < nT * DCEGM_SIM_NMOM; k += kThreads) Sb.moments[(size_t)m * nT * DCEGM_SIM_NMOM + k] = mom[k] / dn;
    for (int k = tid; k < nT + 1; k += kThreads) Sb.ret_hist[(size_t)m * (nT + 1) + k] = hist[k];
    if (tid == 0) Sb.sim_status[m] = s_status;
}

cudaError_t launch_sim(const DeviceBatch& B0, const SimBatch& Sb, int first, int count, int max_smem_optin, cudaStream_t stream) {
    DeviceBatch B = B0;
    B.first = first;
    const bool gl = sim_smem_bytes(B.lay, false) > (size_t)max_smem_optin || B.lay.cap >= 2 * kSearchTop;
    const size_t smem = sim_smem_bytes(B.lay, gl);
    cudaError_t e = cudaSuccess;
#define DC_SIM_LAUNCH(...)                                                                                          \
    do {                                                                                                            \
        e = cudaFuncSetAttribute(sim_kernel<__VA_ARGS__>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  \
        if (e != cudaSuccess) return e;                                                                             \
        sim_kernel<__VA_ARGS__><<<count, kThreads, smem, stream>>>(B, Sb);                                          \
    } while (0)
    const bool small = B.lay.cap < 256;  // na <= ~170: eight search steps instead of eleven
    if (gl) {  // lines too long for shared memory: read in place, 16-step searches (cap < 65536)
        if (B.lay.model == DCEGM_MODEL_F) DC_SIM_LAUNCH(DCEGM_MODEL_F, MemGlobal::kTop, true); else DC_SIM_LAUNCH(DCEGM_MODEL_G, MemGlobal::kTop, true);
    } else if (B.lay.model == DCEGM_MODEL_F) {
        if (small) DC_SIM_LAUNCH(DCEGM_MODEL_F, 128); else DC_SIM_LAUNCH(DCEGM_MODEL_F, kSearchTop);
    } else {
        if (small) DC_SIM_LAUNCH(DCEGM_MODEL_G, 128); else DC_SIM_LAUNCH(DCEGM_MODEL_G, kSearchTop);
    }
#undef DC_SIM_LAUNCH
    return cudaGetLastError();
}

}  // namespace dcegm
