// compact_lines.cu — lossless compact encoding of the stored polylines for the device-to-host transfer.
//
// A stored line is a Vector{Point} (line.jl:1-4, 16 B per point), but almost every abscissa is redundant:
//   * policy lines are (avec + c0, c0) (dc_algo.jl:65): x = a[i] + y for the savings grid point i the point came from;
//   * the value line of the same (period, state, choice) has the abscissae of its policy line (dc_algo.jl:78:
//     same avec + c0; the envelope keeps or drops points of both lines together, dc_algo.jl:482, and the inserted
//     kink pair sits at the intersection's x, dc_algo.jl:486-492).
// The encoder ships the ordinates and, instead of the abscissae, a few RUNS that say where each stretch of
// abscissae comes from; only points for which neither rule reproduces the stored x BIT FOR BIT (left kink points,
// bend-back points of the value line, ...) carry an explicit x.  The rules are verified per point on the device, so
// the encoding is lossless whatever the solver produced (non-monotone or non-finite lines included).
//
// The ordinates themselves are smooth along a line, so they travel as the SECOND DIFFERENCES of their 64-bit patterns
// (integer arithmetic modulo 2^64: exactly invertible), packed in blocks of 32 values with one byte width per block.
//
// Record of one line slot, in 8-byte words (include/dcegm.h, "compact lines"):
//     head | ysec[yw] | run[nrun] | xs[nx]          head = yw (words of the ordinate section)
//     ysec = I_0 | I_1 - I_0 | width bytes w_b of the ceil((n-2)/32) blocks, padded to 8 | payload, padded to 8
//       I_k = bit pattern of y_k; block b holds R_k = I_k - 2 I_(k-1) + I_(k-2) for k = 2 + 32 b .. as little-endian
//       two's-complement integers of w_b bytes (w_b = 0: all zero)
//     run = pos | src << 16 | kind << 32 : points pos .. (next run's pos or n) - 1 take their abscissae from
//       kind 0 (GRID)     x[pos + j] = a[src + j] + y[pos + j]          (one IEEE addition, as the kernels did)
//       kind 1 (REF)      x[pos + j] = x of point src + j of the policy line of the same (model, it, iy, id)
//       kind 2 (EXPLICIT) x[pos + j] = the next unread word of xs
// Two passes per chunk: compact_count_kernel (per-point codes, sizes) -> chunk_offsets_kernel (prefix offsets,
// solver_misc.cu) -> compact_write_kernel (streams lines + codes into the records).  One CTA per (policy, value) slot pair.
#include "device_common.cuh"

namespace dcegm {

namespace {

constexpr int kCT = 256;  // threads per CTA of the two kernels
constexpr unsigned kKindGrid = 0, kKindRef = 1, kKindExplicit = 2;
constexpr unsigned kNoCode = 0xffffffffu;

__device__ __forceinline__ bool same_bits(double a, double b) { return __double_as_longlong(a) == __double_as_longlong(b); }

// number of entries <= q in the ascending array x[0], x[stride], ... (n entries); 0 for NaN
__device__ __forceinline__ int count_le_strided(const double* __restrict__ x, int stride, int n, double q) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (x[(size_t)mid * stride] <= q)
            lo = mid + 1;
        else
            hi = mid;
    }
    return lo;
}

// kind << 16 | src of one point.  Policy points: the grid index is first guessed arithmetically (uniform grids:
// three independent probes), then searched (any ascending grid).  Value points: exact match of the abscissa in the
// policy line, searched in the shared-memory copy xs_c of its abscissae when it was staged (nc <= kStageCap),
// else in global memory.
constexpr int kStageCap = 1024;
__device__ __forceinline__ unsigned classify(double2 pt, bool value_line, const double* __restrict__ a, int na, double a0,
                                             double inv_da, const double2* __restrict__ Lc, int nc,
                                             const double* __restrict__ xs_c) {
    if (!value_line) {
        const double d = pt.x - pt.y;
        const double t = (d - a0) * inv_da;
        if (t > -1.0 && t < (double)na) {  // false for NaN
            const int g = (int)(t + 0.5);
#pragma unroll
            for (int o = 0; o < 3; o++) {
                const int i = g + (o == 0 ? 0 : o == 1 ? -1 : 1);
                if (i >= 0 && i < na && same_bits(a[i] + pt.y, pt.x)) return (kKindGrid << 16) | (unsigned)i;
            }
        }
        const int cnt = count_le_strided(a, 1, na, d);
#pragma unroll
        for (int o = 0; o < 4; o++) {
            const int i = cnt - 1 + (o == 0 ? 0 : o == 1 ? 1 : o == 2 ? -1 : 2);
            if (i >= 0 && i < na && same_bits(a[i] + pt.y, pt.x)) return (kKindGrid << 16) | (unsigned)i;
        }
    } else if (xs_c) {
        const int j = count_le_strided(xs_c, 1, nc, pt.x) - 1;
        if (j >= 0 && same_bits(xs_c[j], pt.x)) return (kKindRef << 16) | (unsigned)j;
    } else {
        const int j = count_le_strided(&Lc[0].x, 2, nc, pt.x) - 1;
        if (j >= 0 && same_bits(Lc[j].x, pt.x)) return (kKindRef << 16) | (unsigned)j;
    }
    return kKindExplicit << 16;
}

__device__ __forceinline__ bool starts_run(unsigned prev, unsigned cur) {
    if (prev == kNoCode || (prev >> 16) != (cur >> 16)) return true;
    return (cur >> 16) != kKindExplicit && (cur & 0xffffu) != (prev & 0xffffu) + 1;
}

struct ClassifyCtx {
    const double* a;
    int na;
    double a0, inv_da;
    const double2* Lc;
    int nc;
    const double* xs_c;  // shared-memory copy of the policy line's abscissae or null
};

// One line: run boundaries and (WRITE) the record.  All kCT threads call it.  The count pass (WRITE = false)
// classifies every point and keeps the codes in `codes` (global scratch, one word per point); the write pass reads
// them back.  force_explicit: the count pass found the run form longer than one explicit run (n + 1 + n words),
// which is then what is written.
template <bool WRITE>
__device__ void encode_line(const double2* __restrict__ L, int n, bool value_line, const ClassifyCtx& cx, unsigned* __restrict__ codes,
                            bool force_explicit, double* __restrict__ rec_runs, double* __restrict__ rec_xs, int* nrun_out,
                            int* nx_out) {
    __shared__ unsigned s_code[kCT + 1];
    __shared__ int s_warp[kCT / 32];
    __shared__ int s_tot;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int run_base = 0, x_base = 0;
    unsigned carry = kNoCode;
    for (int k0 = 0; k0 < n; k0 += kCT) {
        const int k = k0 + tid;
        double2 pt = make_double2(0.0, 0.0);
        unsigned code = kNoCode;
        if (k < n) {
            pt = L[k];
            if (WRITE) {
                code = force_explicit ? (kKindExplicit << 16) : codes[k];
            } else {
                code = classify(pt, value_line, cx.a, cx.na, cx.a0, cx.inv_da, cx.Lc, cx.nc, cx.xs_c);
                codes[k] = code;
            }
        }
        s_code[tid + 1] = code;
        if (tid == 0) s_code[0] = carry;
        __syncthreads();
        const bool flag = k < n && starts_run(s_code[tid], code);
        const bool isx = k < n && (code >> 16) == kKindExplicit;
        // exclusive block scan of (flag, isx) packed into one int (n < 65536: no carry between the halves)
        const int v = (flag ? 1 : 0) | (isx ? 1 << 16 : 0);
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[wid] = inc;
        const unsigned last_code = s_code[(n - k0 < kCT ? n - k0 : kCT)];
        __syncthreads();
        if (tid == 0) {
            int run = 0;
            for (int w = 0; w < kCT / 32; w++) {
                const int t = s_warp[w];
                s_warp[w] = run;
                run += t;
            }
            s_tot = run;
        }
        __syncthreads();
        const int excl = s_warp[wid] + inc - v;
        if (WRITE && k < n) {
            if (flag) {
                const unsigned long long e = (unsigned long long)k | ((unsigned long long)(code & 0xffffu) << 16) |
                                             ((unsigned long long)(code >> 16) << 32);
                rec_runs[run_base + (excl & 0xffff)] = __longlong_as_double((long long)e);
            }
            if (isx) rec_xs[x_base + (excl >> 16)] = pt.x;
        }
        run_base += s_tot & 0xffff;
        x_base += s_tot >> 16;
        carry = last_code;
        __syncthreads();
    }
    *nrun_out = run_base;
    *nx_out = x_base;
}

// bytes of the shortest two's-complement form of r (0 for r == 0)
__device__ __forceinline__ int signed_bytes(long long r) {
    if (r == 0) return 0;
    const int bits = 65 - __clzll(r ^ (r >> 63));  // magnitude bits + sign bit
    return (bits + 7) >> 3;
}
__host__ __device__ inline int ysec_head_words(int n) {  // raw words + width bytes, in words
    const int nblk = n > 2 ? (n - 2 + 31) / 32 : 0;
    return (n < 2 ? n : 2) + (nblk + 7) / 8;
}

// The ordinate section of one line.  Count pass (WRITE = false): returns its size in words.  Write pass: fills `sec`
// (8-byte aligned).  Warps and blocks of 32 residuals coincide, so a block's width is one warp reduction.
template <bool WRITE>
__device__ int encode_y(const double2* __restrict__ L, int n, double* __restrict__ sec) {
    __shared__ int s_blk[kCT / 32];
    __shared__ int s_sum;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (n == 0) return 0;
    const int m = n > 2 ? n - 2 : 0;            // residuals
    const int nblk = (m + 31) / 32;
    const int head = ysec_head_words(n);
    unsigned char* wbytes = reinterpret_cast<unsigned char*>(sec + (n < 2 ? n : 2));
    unsigned char* payload = reinterpret_cast<unsigned char*>(sec + head);
    if (WRITE && tid == 0) {
        const long long i0 = __double_as_longlong(L[0].y);
        sec[0] = __longlong_as_double(i0);
        if (n > 1) sec[1] = __longlong_as_double((long long)((unsigned long long)__double_as_longlong(L[1].y) - (unsigned long long)i0));
        for (int b = nblk; b < ((nblk + 7) / 8) * 8; b++) wbytes[b] = 0;  // padding of the width bytes
    }
    int base = 0;  // payload bytes of the tiles done so far (block-uniform)
    for (int j0 = 0; j0 < m; j0 += kCT) {
        const int j = j0 + tid;
        long long r = 0;
        if (j < m) {
            const unsigned long long a = (unsigned long long)__double_as_longlong(L[j + 2].y);
            const unsigned long long b = (unsigned long long)__double_as_longlong(L[j + 1].y);
            const unsigned long long c = (unsigned long long)__double_as_longlong(L[j].y);
            r = (long long)(a - 2ULL * b + c);
        }
        const int w = __reduce_max_sync(0xffffffffu, j < m ? signed_bytes(r) : 0);
        const int cnt = m - (j0 + wid * 32) < 32 ? (m - (j0 + wid * 32) > 0 ? m - (j0 + wid * 32) : 0) : 32;
        if (lane == 0) s_blk[wid] = w * cnt;
        __syncthreads();
        if (tid == 0) {
            int run = 0;
            for (int q = 0; q < kCT / 32; q++) {
                const int t = s_blk[q];
                s_blk[q] = run;
                run += t;
            }
            s_sum = run;
        }
        __syncthreads();
        if (WRITE && j < m) {
            if (lane == 0) wbytes[(j0 >> 5) + wid] = (unsigned char)w;
            unsigned char* dst = payload + base + s_blk[wid] + lane * w;
            unsigned long long u = (unsigned long long)r;
            for (int i = 0; i < w; i++) {
                dst[i] = (unsigned char)(u & 0xffu);
                u >>= 8;
            }
        }
        base += s_sum;
        __syncthreads();
    }
    if (WRITE && tid < 8) {  // padding of the payload
        const int pad = ((base + 7) & ~7) - base;
        if (tid < pad) payload[base + tid] = 0;
    }
    return head + (base + 7) / 8;
}

}  // namespace

// sizes of the compact records of the slot pairs [first_pair, first_pair + gridDim.x) and the per-point codes:
// aux[slot] = nrun | nx << 16, wsize[slot] = 1 + yw + nrun + nx (words; 0 for an empty slot), codes[slot][cap]
__global__ void __launch_bounds__(kCT) compact_count_kernel(DeviceBatch B, size_t first_pair, int32_t* __restrict__ aux,
                                                            int32_t* __restrict__ wsize, unsigned* __restrict__ codes,
                                                            int32_t* __restrict__ aux_host, int32_t* __restrict__ lens_host) {
    __shared__ double s_xc[kStageCap];
    const size_t q = first_pair + blockIdx.x;
    const size_t sc = 2 * q, sv = sc + 1;
    const int m = (int)(sc / ((size_t)B.lay.nT * B.lay.nstate * 4));
    const DevModel& p = B.models[m];
    const int cap = B.lay.cap;
    int nc = B.lens[sc], nv = B.lens[sv];
    nc = nc < cap ? nc : cap;
    nv = nv < cap ? nv : cap;
    ClassifyCtx cx;
    cx.a = B.grids + (size_t)p.grid * p.na;
    cx.na = p.na;
    cx.a0 = cx.a[0];
    cx.inv_da = p.na > 1 ? 1.0 / (cx.a[1] - cx.a[0]) : 0.0;
    cx.Lc = B.lines + sc * (size_t)cap;
    cx.nc = nc;
    cx.xs_c = nullptr;
    const double2* Lv = B.lines + sv * (size_t)cap;
    if (nc <= kStageCap && nv > 0) {
        for (int k = threadIdx.x; k < nc; k += kCT) s_xc[k] = cx.Lc[k].x;
        cx.xs_c = s_xc;
        __syncthreads();
    }
    for (int which = 0; which < 2; which++) {
        const int n = which ? nv : nc;
        const size_t s = (which ? sv : sc) - 2 * first_pair;
        int nrun, nx;
        encode_line<false>(which ? Lv : cx.Lc, n, which == 1, cx, codes + s * (size_t)cap, false, nullptr, nullptr, &nrun, &nx);
        if (nrun + nx > n + 1) {  // one explicit run is shorter
            nrun = 1;
            nx = n;
        }
        const int yw = encode_y<false>(which ? Lv : cx.Lc, n, nullptr);
        if (threadIdx.x == 0) {
            aux[s] = nrun | (nx << 16);
            wsize[s] = n > 0 ? 1 + yw + nrun + nx : 0;
            if (aux_host) {  // host-mapped copies: no small transfer queues behind the bulk device-to-host copies
                aux_host[s] = nrun | (nx << 16);
                lens_host[s] = n;
            }
        }
    }
}

// writes the records at the word offsets off[] (prefix sums of wsize over the chunk's slots)
__global__ void __launch_bounds__(kCT) compact_write_kernel(DeviceBatch B, size_t first_pair, const int32_t* __restrict__ aux,
                                                            const unsigned* __restrict__ codes, const long long* __restrict__ off,
                                                            double* __restrict__ dst) {
    const size_t q = first_pair + blockIdx.x;
    const size_t sc = 2 * q, sv = sc + 1;
    const int cap = B.lay.cap;
    int nc = B.lens[sc], nv = B.lens[sv];
    nc = nc < cap ? nc : cap;
    nv = nv < cap ? nv : cap;
    ClassifyCtx cx{};
    const double2* Lc = B.lines + sc * (size_t)cap;
    const double2* Lv = B.lines + sv * (size_t)cap;
    for (int which = 0; which < 2; which++) {
        const int n = which ? nv : nc;
        const size_t s = (which ? sv : sc) - 2 * first_pair;
        const int ax = aux[s];
        const int nrun = ax & 0xffff, nx = (ax >> 16) & 0xffff;
        if (n == 0) continue;  // block-uniform
        double* rec = dst + off[s];
        const int yw = encode_y<true>(which ? Lv : Lc, n, rec + 1);
        if (threadIdx.x == 0) rec[0] = __longlong_as_double((long long)yw);
        int r, x;
        encode_line<true>(which ? Lv : Lc, n, which == 1, cx, const_cast<unsigned*>(codes) + s * (size_t)cap,
                          nrun == 1 && nx == n, rec + 1 + yw, rec + 1 + yw + nrun, &r, &x);
    }
}

cudaError_t launch_compact_count(const DeviceBatch& B, int first, int count, int32_t* aux, int32_t* wsize, unsigned* codes,
                                 int32_t* aux_host, int32_t* lens_host, cudaStream_t stream) {
    const size_t pairs_per = (size_t)B.lay.nT * B.lay.nstate * 2;
    if (count <= 0) return cudaSuccess;
    compact_count_kernel<<<(unsigned)(pairs_per * count), kCT, 0, stream>>>(B, pairs_per * first, aux, wsize, codes, aux_host, lens_host);
    return cudaGetLastError();
}
cudaError_t launch_compact_write(const DeviceBatch& B, int first, int count, const int32_t* aux, const unsigned* codes,
                                 const long long* off, double* dst, cudaStream_t stream) {
    const size_t pairs_per = (size_t)B.lay.nT * B.lay.nstate * 2;
    if (count <= 0) return cudaSuccess;
    compact_write_kernel<<<(unsigned)(pairs_per * count), kCT, 0, stream>>>(B, pairs_per * first, aux, codes, off, dst);
    return cudaGetLastError();
}

}  // namespace dcegm
