// solver_misc.cu — minimal consumption-savings EGM kernel and the line packing kernel.
#include <math_constants.h>

#include "device_common.cuh"

namespace dcegm {

// minimal_EGM (src/minimal_EGM.jl:3-43), minimal_EGM_borrow (:49-89) and minimal_EGM_bequest (:92-159):
// log-utility Euler equation, no discrete choice.  One CTA per parameterisation; the whole 24-step
// recursion runs out of shared memory.  Output: kind C lines (m[it], c[it]) for id 0; the bequest
// variant also keeps the value function as kind V lines (m[it], v[it]).
__global__ void __launch_bounds__(kThreads) solve_minimal_kernel(DeviceBatch B) {
    const int nthr = blockDim.x;  // one thread per grid point, at most kThreads (coarse grids get a narrower CTA)
    extern __shared__ __align__(16) unsigned char smem[];
    const int m = B.first + blockIdx.x, tid = threadIdx.x;
    __shared__ DevModel s_p;
    __shared__ int s_status;
    __shared__ double s_ev0;
    if (tid == 0) {
        s_p = B.models[m];
        s_status = 0;
    }
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    const bool borrow = B.lay.model == DCEGM_MODEL_MINIMAL_BORROW, beq = B.lay.model == DCEGM_MODEL_MINIMAL_BEQUEST;
    double* buf[2][3];  // [set][m | c | v]
    for (int s = 0; s < 2; s++)
        for (int a = 0; a < 3; a++) buf[s][a] = (double*)smem + (size_t)(s * 3 + a) * (na + 1);
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny;
    const double* __restrict__ ey = B.off + (size_t)m * ny;  // 0.0 + exp(yvec)
    const double m_lo = borrow ? p.a_low : 0.0;               // minimal_EGM.jl:10, 56, 83
    int cur = 0, nnext = 2;
    if (beq) {  // minimal_EGM.jl:102-104: terminal lines on the whole savings grid
        nnext = na;
        for (int k = tid; k < na; k += nthr) {
            buf[0][0][k] = avec[k];
            buf[0][1][k] = avec[k];
            buf[0][2][k] = p.nu > 0 ? p.nu * u_plain(avec[k] + p.bbar, p) : 0.0;
        }
    } else if (tid == 0) {  // minimal_EGM.jl:10-11, 56-57
        buf[0][0][0] = m_lo; buf[0][1][0] = 0.0;
        buf[0][0][1] = p.a_high; buf[0][1][1] = p.a_high;
    }
    __syncthreads();
    {
        const size_t slot = line_slot(B.lay, m, nT - 1, 0, 0, DCEGM_KIND_C);
        if (tid == 0) B.lens[slot] = nnext;
        if (B.lines)
            for (int k = tid; k < nnext; k += nthr) B.lines[slot * (size_t)B.lay.cap + k] = make_double2(buf[0][0][k], buf[0][1][k]);
        if (beq) {
            const size_t sv = line_slot(B.lay, m, nT - 1, 0, 0, DCEGM_KIND_V);
            if (tid == 0) B.lens[sv] = nnext;
            if (B.lines)
                for (int k = tid; k < nnext; k += nthr) B.lines[sv * (size_t)B.lay.cap + k] = make_double2(buf[0][0][k], buf[0][2][k]);
        }
    }
    for (int t = nT - 2; t >= 0; t--) {
        const double* nx = buf[cur][0];
        const double* nyv = buf[cur][1];
        const double* nv = buf[cur][2];
        double* cx = buf[cur ^ 1][0];
        double* cy = buf[cur ^ 1][1];
        double* cv = buf[cur ^ 1][2];
        const bool beq_period = beq && t == nT - 2 && p.nu > 0;  // minimal_EGM.jl:124-133
        for (int ia = tid; ia < na; ia += nthr) {
            const double Ra = p.R * avec[ia];
            double acc = 0.0, accv = 0.0;
            for (int iy = 0; iy < ny; iy++) {
                const double w1 = ey[iy] + Ra;                 // minimal_EGM.jl:21
                double c1 = interp_y(nx, nyv, nnext, w1);      // minimal_EGM.jl:25
                if (c1 < 0) c1 = p.cfloor;                     // minimal_EGM.jl:26
                const double mu = beq_period ? 1.0 / (c1 + p.bbar) : 1.0 / c1;
                acc = iy == 0 ? wgt[iy] * mu : acc + wgt[iy] * mu;  // minimal_EGM.jl:27
                if (beq) {                                     // minimal_EGM.jl:150-151
                    const double v1 = interp_y(nx, nv, nnext, w1);
                    accv = iy == 0 ? wgt[iy] * v1 : accv + wgt[iy] * v1;
                }
            }
            const double rhs = (beq_period ? p.betaRnu : p.betaR) * acc;  // minimal_EGM.jl:28, 132
            const double c0 = 1.0 / rhs;                       // minimal_EGM.jl:29
            cx[1 + ia] = avec[ia] + c0;                        // minimal_EGM.jl:30
            cy[1 + ia] = c0;
            if (beq) {
                cv[1 + ia] = u_plain(c0, p) + p.beta * accv;   // minimal_EGM.jl:152
                if (ia == 0) s_ev0 = accv;
            }
        }
        __syncthreads();
        if (tid == 0) {  // minimal_EGM.jl:36-37, 82-83, 146, 152
            cx[0] = m_lo;
            cy[0] = 0.0;
            if (beq) cv[0] = s_ev0;
        }
        __syncthreads();
        int flag = 0;
        for (int k = tid; k < na; k += nthr)
            flag |= knot_pair_flag(cx[k], cx[k + 1], t == 0);
        if (flag) atomicOr(&s_status, flag);
        const size_t slot = line_slot(B.lay, m, t, 0, 0, DCEGM_KIND_C);
        if (tid == 0) B.lens[slot] = na + 1;
        if (B.lines)
            for (int k = tid; k < na + 1; k += nthr) B.lines[slot * (size_t)B.lay.cap + k] = make_double2(cx[k], cy[k]);
        if (beq) {
            const size_t sv = line_slot(B.lay, m, t, 0, 0, DCEGM_KIND_V);
            if (tid == 0) B.lens[sv] = na + 1;
            if (B.lines)
                for (int k = tid; k < na + 1; k += nthr) B.lines[sv * (size_t)B.lay.cap + k] = make_double2(cx[k], cv[k]);
        }
        nnext = na + 1;
        cur ^= 1;
        __syncthreads();
    }
    if (tid == 0) {
        B.status[m] = s_status;
        dcegm_summary sm;
        memset(&sm, 0, sizeof sm);
        sm.status = s_status;
        sm.max_len = na + 1;
        sm.tot_points = ((long long)(nT - 1) * (na + 1) + (beq ? na : 2)) * (beq ? 2 : 1);
        const double fr[DCEGM_SUMMARY_NPROBE] = {0.02, 0.05, 0.1, 0.2, 0.3, 0.5, 0.7, 0.9};
        double cs = 0.0;
        for (int k = 0; k < nnext; k++) cs += buf[cur][1][k];
        for (int q = 0; q < DCEGM_SUMMARY_NPROBE; q++) sm.c1_probe[q] = interp_y(buf[cur][0], buf[cur][1], nnext, fr[q] * p.a_high);
        sm.checksum_c = cs;
        B.summaries[m] = sm;
    }
}

cudaError_t launch_solve_minimal(const DeviceBatch& B0, int first, int count, cudaStream_t stream) {
    DeviceBatch B = B0;
    B.first = first;
    const size_t smem = (size_t)6 * (B.lay.na + 1) * sizeof(double);
    if (smem > 48 * 1024) {  // per device, set on every launch (several handles / GPUs per process)
        cudaError_t e = cudaFuncSetAttribute(solve_minimal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    int nthr = (B.lay.na + 1 + 31) / 32 * 32;  // a thread per grid point
    nthr = nthr > kThreads ? kThreads : nthr;
    solve_minimal_kernel<<<count, nthr, smem, stream>>>(B);
    return cudaGetLastError();
}

// dense packing of the stored polylines: one CTA per line slot copies len points to dst+offset
__global__ void pack_lines_kernel(const double2* __restrict__ lines, const long long* __restrict__ offsets, int cap,
                                  size_t first_slot, double2* __restrict__ dst) {
    const size_t s = blockIdx.x;
    const long long o0 = offsets[s], o1 = offsets[s + 1];
    const double2* src = lines + (first_slot + s) * (size_t)cap;
    for (long long k = threadIdx.x; k < o1 - o0; k += blockDim.x) dst[o0 + k] = src[k];
}

// exclusive prefix sum of min(lens, cap) over the nsl line slots of one chunk -> off[0..nsl] (device, read by
// pack_lines_kernel) and the same numbers into host-mapped memory: the host learns the chunk's size and the
// caller's offsets without a transfer that would queue behind the bulk device-to-host copies
__global__ void __launch_bounds__(1024) chunk_offsets_kernel(const int32_t* __restrict__ lens, int nsl, int cap,
                                                             long long* __restrict__ off, long long* __restrict__ off_host) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nsl; b0 += 1024) {
        const int k = b0 + tid;
        long long v = 0;
        if (k < nsl) {
            const int n = lens[k];
            v = n < cap ? n : cap;
        }
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            long long w = s_warp[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;  // exclusive over warps
        }
        __syncthreads();
        const long long carry = s_carry;
        const long long excl = carry + s_warp[wid] + inc - v;
        if (k < nsl) {
            off[k] = excl;
            off_host[k] = excl;
        }
        __syncthreads();
        if (tid == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (tid == 0) {
        off[nsl] = s_carry;
        off_host[nsl] = s_carry;
    }
}

cudaError_t launch_chunk_offsets(const int32_t* lens, size_t nsl, int cap, long long* d_off, long long* h_off_mapped,
                                 cudaStream_t stream) {
    chunk_offsets_kernel<<<1, 1024, 0, stream>>>(lens, (int)nsl, cap, d_off, h_off_mapped);
    return cudaGetLastError();
}

cudaError_t launch_pack_lines(const DeviceBatch& B, int first, int count, const long long* d_offsets, double2* d_dst,
                              cudaStream_t stream) {
    const size_t per = (size_t)B.lay.nT * B.lay.nstate * 4;
    const size_t nsl = per * count;
    if (nsl == 0) return cudaSuccess;
    pack_lines_kernel<<<(unsigned)nsl, 128, 0, stream>>>(B.lines, d_offsets, B.lay.cap, per * first, d_dst);
    return cudaGetLastError();
}

// One fixed-size record per model for the single all-gather of a multi-GPU sweep (include/dcegm.h:
// dcegm_sweep_record_bytes): [dcegm_summary | int32 sim_status | int32 nT | double moments[nT_pad][DCEGM_SIM_NMOM]].
// Moments are NaN when no simulation ran, beyond nT, and for models whose solution the simulator must not be
// trusted on: any fatal status bit, or a first-period line that is not strictly increasing (sim interpolates period 1).
__global__ void sweep_records_kernel(const dcegm_summary* __restrict__ summ, const double* __restrict__ mom,
                                     const int32_t* __restrict__ sim_status, int nT, int nT_pad, size_t rec_bytes,
                                     unsigned char* __restrict__ dst) {
    const int m = blockIdx.x;
    unsigned char* r = dst + (size_t)m * rec_bytes;
    const dcegm_summary sm = summ[m];
    if (threadIdx.x == 0) {
        *reinterpret_cast<dcegm_summary*>(r) = sm;
        int32_t* tail = reinterpret_cast<int32_t*>(r + sizeof(dcegm_summary));
        tail[0] = mom ? sim_status[m] : -1;
        tail[1] = nT;
    }
    double* out = reinterpret_cast<double*>(r + sizeof(dcegm_summary) + 8);
    const bool masked = !mom || ((unsigned)sm.status & (DCEGM_ST_FATAL_MASK | DCEGM_ST_NONMONO_FIRST)) != 0;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    for (int k = threadIdx.x; k < nT_pad * DCEGM_SIM_NMOM; k += blockDim.x)
        out[k] = (masked || k >= nT * DCEGM_SIM_NMOM) ? nan : mom[(size_t)m * nT * DCEGM_SIM_NMOM + k];
}
cudaError_t launch_sweep_records(const dcegm_summary* summ, const double* mom, const int32_t* sim_status, int nmodels, int nT,
                                 int nT_pad, size_t rec_bytes, unsigned char* dst, cudaStream_t stream) {
    sweep_records_kernel<<<nmodels, 128, 0, stream>>>(summ, mom, sim_status, nT, nT_pad, rec_bytes, dst);
    return cudaGetLastError();
}

// ---- diagnostics: device arithmetic of dc_math.h against the IEEE operations -------------------
// Operands are drawn (counter-based splitmix64) from the ranges the solver kernels use; which[0]
// counts dc_div_fast(a,b) != a/b, which[1] dc_rcp_fast(b) != 1/b.
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
__global__ void math_check_kernel(unsigned long long n, unsigned long long seed, int emin, int emax,
                                  unsigned long long* which) {
    unsigned long long bad_div = 0, bad_rcp = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long u = mix64(seed + 3 * i), v = mix64(seed + 3 * i + 1), w = mix64(seed + 3 * i + 2);
        const int span = emax - emin + 1;
        const long long ea = emin + (long long)(w % (unsigned)span), eb = emin + (long long)((w >> 20) % (unsigned)span);
        double a = __longlong_as_double(((ea + 1023) << 52) | (long long)(u >> 12));
        double b = __longlong_as_double(((eb + 1023) << 52) | (long long)(v >> 12));
        if (w >> 63) a = -a;
        if ((w >> 62) & 1) b = -b;
        if (((w >> 40) & 1023) == 0) a = 0.0;          // exact hits of a knot
        if (((w >> 50) & 255) == 0) a = b * (double)((u >> 60) + 1);  // exact quotients
        bad_div += dc_div_fast(a, b) != a / b;
        bad_rcp += dc_rcp_fast(b) != 1.0 / b;
    }
    if (bad_div) atomicAdd(&which[0], bad_div);
    if (bad_rcp) atomicAdd(&which[1], bad_rcp);
}

// ---- diagnostics: FP64 issue rate of this device ------------------------------------------------
// Eight independent DFMA chains per thread, 1024 resident threads per SM, no memory traffic: the rate at
// which the FP64 pipe retires warp instructions.  The denominator of the FP64 roofline in bench.py.
__global__ void __launch_bounds__(256) fp64_peak_kernel(int iters, double seed, double* sink) {
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, c = 1e-7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    const double r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (r == 12345.678) sink[0] = r;  // never true: keeps the chains alive
}

}  // namespace dcegm

extern "C" int dcegm_debug_dfma_rate(int device_ordinal, double* dfma_per_s, double* ms) {
    if (!dfma_per_s) return DCEGM_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return DCEGM_E_NODEVICE;
    if (device_ordinal < 0 || device_ordinal >= ndev || cudaSetDevice(device_ordinal) != cudaSuccess) return DCEGM_E_INVALID;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_ordinal);
    double* d = nullptr;
    if (cudaMalloc(&d, 8) != cudaSuccess) return DCEGM_E_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 4096, blocks = sms * 4 * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {  // first repetition warms up
        cudaEventRecord(e0);
        dcegm::fp64_peak_kernel<<<blocks, 256>>>(iters, 1.0 + rep, d);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float t = 0.f;
        cudaEventElapsedTime(&t, e0, e1);
        if (rep > 0 && t < best) best = t;
    }
    const cudaError_t e = cudaGetLastError();
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    if (e != cudaSuccess || best > 1e29f) return DCEGM_E_CUDA;
    *dfma_per_s = (double)blocks * 256.0 * iters * 64.0 / (best * 1e-3);
    if (ms) *ms = best;
    return DCEGM_OK;
}

extern "C" int dcegm_debug_math_check(unsigned long long n, unsigned long long seed, int emin, int emax,
                                      unsigned long long* mismatches2) {
    if (!mismatches2 || emin < -1000 || emax > 1000 || emin > emax) return DCEGM_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return DCEGM_E_NODEVICE;
    unsigned long long* d = nullptr;
    if (cudaMalloc(&d, 16) != cudaSuccess) return DCEGM_E_CUDA;
    cudaMemset(d, 0, 16);
    dcegm::math_check_kernel<<<148 * 8, 256>>>(n, seed, emin, emax, d);
    cudaError_t e = cudaMemcpy(mismatches2, d, 16, cudaMemcpyDeviceToHost);
    cudaFree(d);
    return e == cudaSuccess ? DCEGM_OK : DCEGM_E_CUDA;
}
