// no_alloc.cu — dcegm_pack_lines_noalloc: the dense packing of stored polylines (dcegm_fetch_packed: prefix offsets
// of the line lengths, then one copy per line slot) for callers that must not allocate — code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc.  The existing path keeps its prefix offsets in a device
// buffer grown with cudaMalloc (dcegm_abi.cu: pack_off); this variant keeps them in a file-scope static device array
// of fixed size.  No counterpart in the reference.
//
// The array is part of the library image: every GPU that loads the library holds its 8 MiB until the process
// exits.  All calls on a device share it and it keeps the last call's contents; the one word that must start at zero
// (the count of slots that did not fit) is cleared on the caller's stream at the start of every call.  Two calls whose
// GPU work overlaps (different streams, or different host threads without ordering) race on it.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/dcegm.h"

namespace dcegm {

namespace {

constexpr size_t kCtrlBytes = 256;  // control block in front of the offsets: [0] total points, [1] slots that did not fit
constexpr size_t kScratchBytes = kCtrlBytes + (((size_t)DCEGM_NOALLOC_MAX_SLOTS + 1) * sizeof(long long) + 255) / 256 * 256;

// In host code this name is a host-side stand-in, not a device address: the launches below get the address that
// cudaGetSymbolAddress returns for the current device (scratch_address()).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// exclusive prefix sum of min(lens, cap) over nsl slots -> off[0..nsl] (scratch) and, when asked for, off_out (caller's
// device memory); ctrl[0] = total.  One CTA: the scan is a few microseconds next to the copy it precedes.
__global__ void __launch_bounds__(1024) noalloc_offsets_kernel(const int32_t* __restrict__ lens, long long nsl, int cap,
                                                               long long* __restrict__ off, long long* __restrict__ off_out,
                                                               long long* __restrict__ ctrl) {
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (long long b0 = 0; b0 < nsl; b0 += 1024) {
        const long long k = b0 + tid;
        long long v = 0;
        if (k < nsl) {
            const int n = lens[k];
            v = n < 0 ? 0 : n < cap ? n : cap;
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
            const long long w = s_warp[lane];
            long long wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        const long long excl = s_carry + s_warp[wid] + inc - v;
        if (k < nsl) {
            off[k] = excl;
            if (off_out) off_out[k] = excl;
        }
        __syncthreads();
        if (tid == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (tid == 0) {
        off[nsl] = s_carry;
        if (off_out) off_out[nsl] = s_carry;
        ctrl[0] = s_carry;
    }
}

// one CTA per line slot: copies its points behind those of the slots before it; a slot that would end beyond
// cap_points is left out and counted in ctrl[1]
__global__ void noalloc_pack_kernel(const double2* __restrict__ lines, const long long* __restrict__ off, int cap,
                                    double2* __restrict__ dst, long long cap_points, long long* __restrict__ ctrl) {
    const size_t s = blockIdx.x;
    const long long o0 = off[s], o1 = off[s + 1];
    if (o1 > cap_points) {
        if (threadIdx.x == 0 && o1 > o0) atomicAdd((unsigned long long*)&ctrl[1], 1ULL);
        return;
    }
    const double2* src = lines + s * (size_t)cap;
    for (long long k = threadIdx.x; k < o1 - o0; k += blockDim.x) dst[o0 + k] = src[k];
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices];  // device address of g_noalloc_scratch per device, fetched once

int say(char* buf, size_t cap, int rc, const char* fmt, ...) {
    if (buf && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, cap, fmt, ap);
        va_end(ap);
    }
    return rc;
}

// device address of the static array on device `dev`, which must be the current device
cudaError_t scratch_address(int dev, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    return cudaSuccess;
}

}  // namespace

}  // namespace dcegm

extern "C" int dcegm_pack_lines_noalloc(int device_ordinal, const double* lines_dev, const int32_t* lens_dev, int64_t nslots,
                                        int cap, double* xy_dev, size_t cap_points, int64_t* offsets_dev, int64_t* npoints,
                                        void* stream, char* errbuf, size_t errbuf_bytes) {
    using namespace dcegm;
    if (errbuf && errbuf_bytes) errbuf[0] = 0;
    if (!lines_dev || !lens_dev || !xy_dev)
        return say(errbuf, errbuf_bytes, DCEGM_E_INVALID, "dcegm_pack_lines_noalloc: lines_dev / lens_dev / xy_dev is null");
    if (nslots < 0 || cap < 1) return say(errbuf, errbuf_bytes, DCEGM_E_INVALID, "dcegm_pack_lines_noalloc: nslots < 0 or cap < 1");
    if (nslots > (int64_t)DCEGM_NOALLOC_MAX_SLOTS)
        return say(errbuf, errbuf_bytes, DCEGM_E_INVALID,
                   "dcegm_pack_lines_noalloc: %lld line slots exceed the static scratch (DCEGM_NOALLOC_MAX_SLOTS = %d); "
                   "pack in pieces or use dcegm_fetch_packed",
                   (long long)nslots, (int)DCEGM_NOALLOC_MAX_SLOTS);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return say(errbuf, errbuf_bytes, DCEGM_E_NODEVICE, "dcegm_pack_lines_noalloc: no CUDA device");
    if (device_ordinal < 0 || device_ordinal >= ndev || device_ordinal >= kMaxDevices)
        return say(errbuf, errbuf_bytes, DCEGM_E_INVALID, "dcegm_pack_lines_noalloc: device %d out of range", device_ordinal);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device_ordinal) e = cudaSetDevice(device_ordinal);
    if (e != cudaSuccess) return say(errbuf, errbuf_bytes, DCEGM_E_CUDA, "dcegm_pack_lines_noalloc: %s", cudaGetErrorString(e));
    int rc = DCEGM_OK;
    do {
        unsigned char* scratch = nullptr;
        if ((e = scratch_address(device_ordinal, &scratch)) != cudaSuccess) {
            rc = say(errbuf, errbuf_bytes, DCEGM_E_CUDA, "dcegm_pack_lines_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            break;
        }
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        long long* ctrl = reinterpret_cast<long long*>(scratch);
        long long* off = reinterpret_cast<long long*>(scratch + kCtrlBytes);
        // the array keeps what the last call left: the overflow count (and the total) must start at zero
        if ((e = cudaMemsetAsync(ctrl, 0, kCtrlBytes, st)) != cudaSuccess) {
            rc = say(errbuf, errbuf_bytes, DCEGM_E_CUDA, "dcegm_pack_lines_noalloc: clearing the control block: %s", cudaGetErrorString(e));
            break;
        }
        static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
        noalloc_offsets_kernel<<<1, 1024, 0, st>>>(lens_dev, (long long)nslots, cap, off, reinterpret_cast<long long*>(offsets_dev), ctrl);
        if (nslots > 0)
            noalloc_pack_kernel<<<(unsigned)nslots, 128, 0, st>>>(reinterpret_cast<const double2*>(lines_dev), off, cap,
                                                                  reinterpret_cast<double2*>(xy_dev), (long long)cap_points, ctrl);
        if ((e = cudaGetLastError()) != cudaSuccess) {
            rc = say(errbuf, errbuf_bytes, DCEGM_E_CUDA, "dcegm_pack_lines_noalloc: launch: %s", cudaGetErrorString(e));
            break;
        }
        if (npoints) {  // the caller wants the total on the host: wait for the stream and read the control block
            long long host_ctrl[2] = {0, 0};
            e = cudaMemcpyAsync(host_ctrl, ctrl, sizeof host_ctrl, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) {
                rc = say(errbuf, errbuf_bytes, DCEGM_E_CUDA, "dcegm_pack_lines_noalloc: %s", cudaGetErrorString(e));
                break;
            }
            *npoints = host_ctrl[0];
            if (host_ctrl[1] != 0)
                rc = say(errbuf, errbuf_bytes, DCEGM_E_CAPACITY,
                         "dcegm_pack_lines_noalloc: destination too small: %lld points, room for %zu (%lld slots left out)",
                         host_ctrl[0], cap_points, host_ctrl[1]);
        }
    } while (0);
    if (prev >= 0 && prev != device_ordinal) cudaSetDevice(prev);
    return rc;
}
