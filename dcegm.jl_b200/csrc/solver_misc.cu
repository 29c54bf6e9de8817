// solver_misc.cu — minimal consumption-savings EGM kernel and the line packing kernel.
#include <math_constants.h>

#include "device_common.cuh"

namespace dcegm {

// minimal_EGM (src/minimal_EGM.jl:3-43): log utility, no discrete choice, no value function.
// One CTA per parameterisation; the whole 24-step recursion runs out of shared memory.
// Output: kind C lines of (m[it], c[it]) for id 0.
__global__ void __launch_bounds__(kThreads) solve_minimal_kernel(DeviceBatch B) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int m = blockIdx.x, tid = threadIdx.x;
    __shared__ DevModel s_p;
    __shared__ int s_status;
    if (tid == 0) {
        s_p = B.models[m];
        s_status = 0;
    }
    __syncthreads();
    const DevModel& p = s_p;
    const int na = p.na, ny = p.ny, nT = p.nT;
    double* buf[2][2];
    for (int s = 0; s < 2; s++)
        for (int a = 0; a < 2; a++) buf[s][a] = (double*)smem + (size_t)(s * 2 + a) * (na + 1);
    const double* __restrict__ avec = B.grids + (size_t)p.grid * na;
    const double* __restrict__ wgt = B.wgt + (size_t)m * ny;
    const double* __restrict__ ey = B.off + (size_t)m * ny;  // 0.0 + exp(yvec)
    int cur = 0, nnext = 2;
    if (tid == 0) {  // minimal_EGM.jl:10-11
        buf[0][0][0] = 0.0; buf[0][1][0] = 0.0;
        buf[0][0][1] = p.a_high; buf[0][1][1] = p.a_high;
    }
    __syncthreads();
    {
        const size_t slot = line_slot(B.lay, m, nT - 1, 0, 0, DCEGM_KIND_C);
        if (tid == 0) B.lens[slot] = 2;
        if (B.lines && tid < 2) B.lines[slot * (size_t)B.lay.cap + tid] = make_double2(buf[0][0][tid], buf[0][1][tid]);
    }
    for (int t = nT - 2; t >= 0; t--) {
        const double* nx = buf[cur][0];
        const double* nyv = buf[cur][1];
        double* cx = buf[cur ^ 1][0];
        double* cy = buf[cur ^ 1][1];
        for (int ia = tid; ia < na; ia += kThreads) {
            const double Ra = p.R * avec[ia];
            double acc = 0.0;
            for (int iy = 0; iy < ny; iy++) {
                const double w1 = ey[iy] + Ra;                 // minimal_EGM.jl:21
                double c1 = interp_y(nx, nyv, nnext, w1);      // minimal_EGM.jl:25
                if (c1 < 0) c1 = p.cfloor;                     // minimal_EGM.jl:26
                acc = iy == 0 ? wgt[iy] * (1.0 / c1) : acc + wgt[iy] * (1.0 / c1);  // minimal_EGM.jl:27
            }
            const double rhs = p.betaR * acc;                  // minimal_EGM.jl:28
            const double c0 = 1.0 / rhs;                       // minimal_EGM.jl:29
            cx[1 + ia] = avec[ia] + c0;                        // minimal_EGM.jl:30
            cy[1 + ia] = c0;
        }
        if (tid == 0) {  // minimal_EGM.jl:36-37
            cx[0] = 0.0;
            cy[0] = 0.0;
        }
        __syncthreads();
        int flag = 0;
        for (int k = tid; k < na; k += kThreads)
            if (!(cx[k + 1] > cx[k])) flag = DCEGM_ST_NONMONOTONE;
        if (flag) atomicOr(&s_status, flag);
        const size_t slot = line_slot(B.lay, m, t, 0, 0, DCEGM_KIND_C);
        if (tid == 0) B.lens[slot] = na + 1;
        if (B.lines)
            for (int k = tid; k < na + 1; k += kThreads) B.lines[slot * (size_t)B.lay.cap + k] = make_double2(cx[k], cy[k]);
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
        sm.tot_points = (long long)(nT - 1) * (na + 1) + 2;
        const double fr[DCEGM_SUMMARY_NPROBE] = {0.02, 0.05, 0.1, 0.2, 0.3, 0.5, 0.7, 0.9};
        double cs = 0.0;
        for (int k = 0; k < nnext; k++) cs += buf[cur][1][k];
        for (int q = 0; q < DCEGM_SUMMARY_NPROBE; q++) sm.c1_probe[q] = interp_y(buf[cur][0], buf[cur][1], nnext, fr[q] * p.a_high);
        sm.checksum_c = cs;
        B.summaries[m] = sm;
    }
}

cudaError_t launch_solve_minimal(const DeviceBatch& B, cudaStream_t stream) {
    const size_t smem = (size_t)4 * (B.lay.na + 1) * sizeof(double);
    static size_t configured = 48 * 1024;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(solve_minimal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    solve_minimal_kernel<<<B.lay.nmodels, kThreads, smem, stream>>>(B);
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

cudaError_t launch_pack_lines(const DeviceBatch& B, int first, int count, const long long* d_offsets, double2* d_dst,
                              cudaStream_t stream) {
    const size_t per = (size_t)B.lay.nT * B.lay.nstate * 4;
    const size_t nsl = per * count;
    if (nsl == 0) return cudaSuccess;
    pack_lines_kernel<<<(unsigned)nsl, 128, 0, stream>>>(B.lines, d_offsets, B.lay.cap, per * first, d_dst);
    return cudaGetLastError();
}

}  // namespace dcegm
