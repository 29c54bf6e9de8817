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
// The ordinates themselves are smooth along a line, so they travel as SECOND (or third) DIFFERENCES of their 64-bit
// patterns (integer arithmetic modulo 2^64: exactly invertible), packed in blocks of 8 values with one control byte
// per block: the width of the block's residuals in 4-bit units and which of the two differences it holds.
//
// Record of one line slot, in 8-byte words (include/dcegm.h, "compact lines"):
//     head | ysec[yw] | run[nrun] | xs[nx]          head = yw (words of the ordinate section)
//     ysec = I_0 | I_1 - I_0 | control bytes c_b of the ceil(n/8) blocks, padded to 8 | payload, padded to 8
//       I_k = bit pattern of y_k; R2_k = I_k - 2 I_(k-1) + I_(k-2) for k >= 2, R2_0 = R2_1 = 0 (the first two ordinates
//       are the raw words in front); R3_k = R2_k - R2_(k-1).  Block b covers k = 8 b .. 8 b + 7 (cnt = 8, or what is left
//       of the line) and holds R2 (c_b bit 5 clear) or R3 (bit 5 set) as two's-complement integers of w = c_b & 31
//       four-bit units (0: all zero ... 16: raw), whichever is narrower: first cnt x (w >> 1) bytes, the low bytes of
//       residual j little-endian at byte j (w >> 1); then, for odd w, ceil(cnt / 2) bytes with the top four bits of the
//       residuals 2 i (low half) and 2 i + 1 (high half) in byte i
//     run = pos | src << 16 | kind << 32 : points pos .. (next run's pos or n) - 1 take their abscissae from
//       kind 0 (GRID)     x[pos + j] = a[src + j] + y[pos + j]          (one IEEE addition, as the kernels did)
//       kind 1 (REF)      x[pos + j] = x of point src + j of the policy line of the same (model, it, iy, id)
//       kind 2 (EXPLICIT) x[pos + j] = the next unread word of xs
// Two passes per chunk: compact_count_kernel (per-point codes, sizes) -> chunk_offsets_kernel (prefix offsets,
// solver_misc.cu) -> compact_write_kernel (streams lines + codes into the records).  One WARP per (policy, value) slot
// pair, 32 points per step = four blocks of the ordinate section: ballots and shuffles only, no block-wide barrier.
#include "device_common.cuh"

namespace dcegm {

namespace {

constexpr int kCT = 256;  // threads per CTA of the two kernels
constexpr int kPairsPerCta = kCT / 32;
constexpr unsigned kKindGrid = 0, kKindRef = 1, kKindExplicit = 2;
constexpr unsigned kNoCode = 0xffffffffu;
constexpr int32_t kAliasMark = -1;  // aux / yws of a slot without a record of its own (nrun = nx = 65535 cannot occur)

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
constexpr int kStageCap = 640;  // doubles per warp: 8 x 5 KB of shared memory per CTA
__device__ __forceinline__ unsigned classify(double2 pt, bool value_line, const double* __restrict__ a, int na, double a0,
                                             double inv_da, const double2* __restrict__ Lc, int nc,
                                             const double* __restrict__ xs_c, int hint) {
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
        // value points mostly sit at a constant offset from their policy points: try where the previous step ended
        if (hint >= 0 && hint < nc && same_bits(xs_c[hint], pt.x)) return (kKindRef << 16) | (unsigned)hint;
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
    double* stage;       // count pass over the policy line: where that copy is written (same memory), or null
};

// four-bit units of the shortest two's-complement form of r (0 for r == 0, at most 16)
__device__ __forceinline__ int signed_nibbles(long long r) {
    if (r == 0) return 0;
    const int bits = 65 - __clzll(r ^ (r >> 63));  // magnitude bits + sign bit
    return (bits + 3) >> 2;
}
__host__ __device__ inline int ysec_head_words(int n) {  // raw words + control bytes, in words
    const int nblk = (n + 7) / 8;
    return (n < 2 ? n : 2) + (nblk + 7) / 8;
}

struct LineSizes {
    int nrun, nx, yw;
};

// One line, one warp; 32 points per step.  Count pass (WRITE = false): classifies every point, keeps the codes in
// `codes` (global scratch, one word per point) and returns the sizes.  Write pass: reads the codes back and fills the
// record rec = head | ysec[yw] | run[nrun] | xs[nx] (sizes known from the count pass).  force_explicit: the count pass
// found the run form longer than one explicit run (n + 1 + n words), which is then what is written.
template <bool WRITE>
__device__ LineSizes encode_line(const double2* __restrict__ L, int n, bool value_line, const ClassifyCtx& cx,
                                 unsigned* __restrict__ codes, bool force_explicit, LineSizes known, double* __restrict__ rec) {
    const int lane = threadIdx.x & 31;
    LineSizes z = {0, 0, 0};
    if (n == 0) return z;
    const int head = ysec_head_words(n);
    double* sec = rec + 1;
    unsigned char* ctl = WRITE ? reinterpret_cast<unsigned char*>(sec + (n < 2 ? n : 2)) : nullptr;
    unsigned char* payload = WRITE ? reinterpret_cast<unsigned char*>(sec + head) : nullptr;
    double* rec_runs = WRITE ? rec + 1 + known.yw : nullptr;
    double* rec_xs = WRITE ? rec_runs + known.nrun : nullptr;
    int run_base = 0, x_base = 0, pay = 0;
    unsigned carry = kNoCode;
    unsigned long long ym1 = 0, ym2 = 0;  // bit patterns of the two ordinates before this step
    long long r2m1 = 0;                   // second difference of the ordinate before this step
    int ref_delta = 0;                    // count pass, value lines: policy index minus own index at the end of the last step
    // the loads of the next step are issued before this step's work (one warp has nothing else to hide them behind)
    double2 pt_next = lane < n ? L[lane] : make_double2(0.0, 0.0);
    unsigned stash_next = (WRITE && lane < n) ? codes[lane] : 0u;
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int k = k0 + lane;
        const bool in = k < n;
        const double2 pt = pt_next;
        unsigned code = kNoCode;
        const unsigned stash = stash_next;  // the count pass's word of this point: abscissa code | control byte of its block << 24
        if (k + 32 < n) {
            pt_next = L[k + 32];
            if (WRITE) stash_next = codes[k + 32];
        } else {
            pt_next = make_double2(0.0, 0.0);
            stash_next = 0u;
        }
        if (in) {
            if (!WRITE && !value_line && cx.stage) cx.stage[k] = pt.x;  // read by the value line's pass that follows
            if (WRITE) {
                code = force_explicit ? (kKindExplicit << 16) : (stash & 0x3ffffu);
            } else {
                code = classify(pt, value_line, cx.a, cx.na, cx.a0, cx.inv_da, cx.Lc, cx.nc, cx.xs_c, k + ref_delta);
            }
        }
        // ---- abscissa runs ----
        unsigned prev = __shfl_up_sync(0xffffffffu, code, 1);
        if (lane == 0) prev = carry;
        const bool flag = in && starts_run(prev, code);
        const bool isx = in && (code >> 16) == kKindExplicit;
        const unsigned mflag = __ballot_sync(0xffffffffu, flag), mx = __ballot_sync(0xffffffffu, isx);
        const unsigned below = (1u << lane) - 1u;
        if (WRITE && in) {
            if (flag) {
                const unsigned long long e = (unsigned long long)k | ((unsigned long long)(code & 0xffffu) << 16) |
                                             ((unsigned long long)(code >> 16) << 32);
                rec_runs[run_base + __popc(mflag & below)] = __longlong_as_double((long long)e);
            }
            if (isx) rec_xs[x_base + __popc(mx & below)] = pt.x;
        }
        run_base += __popc(mflag);
        x_base += __popc(mx);
        const int last = (n - k0 < 32 ? n - k0 : 32) - 1;
        carry = __shfl_sync(0xffffffffu, code, last);
        if (!WRITE && (carry >> 16) == kKindRef) ref_delta = (int)(carry & 0xffffu) - (k0 + last);
        // ---- ordinates: second / third differences of the bit patterns, four blocks of 8 per step ----
        const unsigned long long iy = (unsigned long long)__double_as_longlong(pt.y);
        unsigned long long p1 = __shfl_up_sync(0xffffffffu, iy, 1), p2 = __shfl_up_sync(0xffffffffu, iy, 2);
        if (lane == 0) {
            p1 = ym1;
            p2 = ym2;
        } else if (lane == 1) {
            p2 = ym1;
        }
        const long long r2 = (in && k >= 2) ? (long long)(iy - 2ULL * p1 + p2) : 0LL;
        long long r2p = __shfl_up_sync(0xffffffffu, r2, 1);
        if (lane == 0) r2p = r2m1;
        const long long r3 = (in && k >= 2) ? (long long)((unsigned long long)r2 - (unsigned long long)r2p) : 0LL;
        const int j = lane & 7, g = lane >> 3;
        int w;
        bool third;
        if (WRITE) {  // chosen by the count pass
            w = (int)(stash >> 24) & 31;
            third = ((stash >> 24) & 32u) != 0;
        } else {
            // maxima of both widths over the aligned groups of 8 lanes: bytewise maximum of the packed pair
            unsigned ww = (unsigned)signed_nibbles(r2) | ((unsigned)signed_nibbles(r3) << 8);
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) ww = __vmaxu4(ww, __shfl_xor_sync(0xffffffffu, ww, o));
            const int w2 = (int)(ww & 0xffu), w3 = (int)(ww >> 8);
            third = w3 < w2;
            w = third ? w3 : w2;
            if (in) codes[k] = code | ((unsigned)(w | (third ? 32 : 0)) << 24);
        }
        const int hb = w >> 1;
        const long long r = third ? r3 : r2;
        int cntg = n - (k0 + 8 * g);
        cntg = cntg < 0 ? 0 : cntg > 8 ? 8 : cntg;
        const int gbytes = cntg * hb + ((w & 1) ? (cntg + 1) >> 1 : 0);  // at most 68
        // the four blocks' sizes packed into one word (byte g = block g), gathered by one OR reduction
        const unsigned gb4 = __reduce_or_sync(0xffffffffu, j == 0 ? (unsigned)gbytes << (8 * g) : 0u);
        const int goff = (int)__vsadu4(gb4 & ((1u << (8 * g)) - 1u), 0u);
        if (WRITE) {
            const int nib = (w & 1) ? (int)(((unsigned long long)r >> (8 * hb)) & 0xfULL) : 0;
            const int nib_next = __shfl_down_sync(0xffffffffu, nib, 1);
            if (in) {
                if (j == 0) ctl[(k0 >> 3) + g] = (unsigned char)(w | (third ? 32 : 0));
                unsigned char* dst = payload + pay + goff;
                unsigned long long u = (unsigned long long)r;
                for (int i = 0; i < hb; i++) {
                    dst[j * hb + i] = (unsigned char)(u & 0xffu);
                    u >>= 8;
                }
                if ((w & 1) && !(j & 1)) dst[cntg * hb + (j >> 1)] = (unsigned char)(nib | ((j + 1 < cntg ? nib_next : 0) << 4));
            }
        }
        pay += (int)__vsadu4(gb4, 0u);
        ym1 = __shfl_sync(0xffffffffu, iy, last);
        ym2 = last >= 1 ? __shfl_sync(0xffffffffu, iy, last - 1) : ym1;  // (a one-point tail only ends a line)
        r2m1 = __shfl_sync(0xffffffffu, r2, last);
    }
    if (WRITE) {
        const int nblk = (n + 7) / 8;
        if (lane == 0) {
            rec[0] = __longlong_as_double((long long)known.yw);
            const unsigned long long i0 = (unsigned long long)__double_as_longlong(L[0].y);
            sec[0] = __longlong_as_double((long long)i0);
            if (n > 1) sec[1] = __longlong_as_double((long long)((unsigned long long)__double_as_longlong(L[1].y) - i0));
        }
        if (lane < 8 && nblk + lane < ((nblk + 7) / 8) * 8) ctl[nblk + lane] = 0;  // padding of the control bytes
        const int pad = ((pay + 7) & ~7) - pay;
        if (lane < pad) payload[pay + lane] = 0;                                   // ... and of the payload
    }
    z.nrun = run_base;
    z.nx = x_base;
    z.yw = head + (pay + 7) / 8;
    return z;
}

// all n points of A and R bit-identical? (one warp; four 32-point steps in flight per vote, leaves at the first
// vote that saw a difference)
__device__ __forceinline__ bool same_line(const double2* __restrict__ A, const double2* __restrict__ R, int n) {
    const int lane = threadIdx.x & 31;
    const ulonglong2* a = reinterpret_cast<const ulonglong2*>(A);
    const ulonglong2* r = reinterpret_cast<const ulonglong2*>(R);
    for (int k0 = 0; k0 < n; k0 += 128) {
        bool ne = false;
        ulonglong2 u[4], v[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int k = k0 + 32 * q + lane;
            if (k < n) {
                u[q] = a[k];
                v[q] = r[k];
            } else {
                u[q] = v[q] = make_ulonglong2(0ULL, 0ULL);
            }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) ne |= u[q].x != v[q].x || u[q].y != v[q].y;
        if (__any_sync(0xffffffffu, ne)) return false;
    }
    return true;
}

}  // namespace

// sizes of the compact records of the slot pairs [first_pair, first_pair + npairs) and the per-point codes:
// aux[slot] = nrun | nx << 16, yws[slot] = words of the ordinate section, wsize[slot] = 1 + yw + nrun + nx (words; 0 for
// an empty slot), codes[slot][cap].
// alias_pairs > 0 (= slot pairs per model): a pair of a later model whose two lines are bit-identical to the same
// (it, iy, id) pair of the range's FIRST model gets no record — aux = yws = kAliasMark, wsize = 0 — and the host points
// its offsets at the first model's records.  In a sweep over (lambda, sigma) every retiree line is such a line: the
// retiree has no income draw (param.jl:212) and no taste-shocked choice (dc_algo.jl:44).
__global__ void __launch_bounds__(kCT) compact_count_kernel(DeviceBatch B, size_t first_pair, size_t npairs, size_t alias_pairs,
                                                            int32_t* __restrict__ aux, int32_t* __restrict__ yws,
                                                            int32_t* __restrict__ wsize, unsigned* __restrict__ codes,
                                                            int32_t* __restrict__ aux_host, int32_t* __restrict__ lens_host) {
    __shared__ double s_xc[kPairsPerCta][kStageCap];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t lq = (size_t)blockIdx.x * kPairsPerCta + wid;
    if (lq >= npairs) return;  // warp-uniform
    const size_t q = first_pair + lq;
    const size_t sc = 2 * q, sv = sc + 1;
    const int m = (int)(sc / ((size_t)B.lay.nT * B.lay.nstate * 4));
    const DevModel& p = B.models[m];
    const int cap = B.lay.cap;
    int nc = B.lens[sc], nv = B.lens[sv];
    nc = nc < cap ? nc : cap;
    nv = nv < cap ? nv : cap;
    if (alias_pairs && lq >= alias_pairs && nc + nv > 0) {
        const size_t rc = 2 * (first_pair + lq % alias_pairs), rv = rc + 1;  // the first model's slots
        int rnc = B.lens[rc], rnv = B.lens[rv];
        rnc = rnc < cap ? rnc : cap;
        rnv = rnv < cap ? rnv : cap;
        if (rnc == nc && rnv == nv && same_line(B.lines + sc * (size_t)cap, B.lines + rc * (size_t)cap, nc) &&
            same_line(B.lines + sv * (size_t)cap, B.lines + rv * (size_t)cap, nv)) {
            if (lane < 2) {
                const size_t s = sc + lane - 2 * first_pair;
                aux[s] = kAliasMark;
                yws[s] = kAliasMark;
                wsize[s] = 0;
                if (aux_host) {
                    aux_host[s] = kAliasMark;
                    lens_host[s] = lane ? nv : nc;
                }
            }
            return;  // warp-uniform
        }
    }
    ClassifyCtx cx;
    cx.a = B.grids + (size_t)p.grid * p.na;
    cx.na = p.na;
    cx.a0 = cx.a[0];
    cx.inv_da = p.na > 1 ? 1.0 / (cx.a[1] - cx.a[0]) : 0.0;
    cx.Lc = B.lines + sc * (size_t)cap;
    cx.nc = nc;
    cx.xs_c = nullptr;
    cx.stage = nullptr;
    const double2* Lv = B.lines + sv * (size_t)cap;
    if (nc <= kStageCap && nv > 0) cx.xs_c = cx.stage = s_xc[wid];  // filled while the policy line is encoded
    for (int which = 0; which < 2; which++) {
        const int n = which ? nv : nc;
        const size_t s = (which ? sv : sc) - 2 * first_pair;
        const LineSizes none = {0, 0, 0};
        LineSizes z = encode_line<false>(which ? Lv : cx.Lc, n, which == 1, cx, codes + s * (size_t)cap, false, none, nullptr);
        if (z.nrun + z.nx > n + 1) {  // one explicit run is shorter
            z.nrun = 1;
            z.nx = n;
        }
        if (lane == 0) {
            aux[s] = z.nrun | (z.nx << 16);
            yws[s] = z.yw;
            wsize[s] = n > 0 ? 1 + z.yw + z.nrun + z.nx : 0;
            if (aux_host) {  // host-mapped copies: no small transfer queues behind the bulk device-to-host copies
                aux_host[s] = z.nrun | (z.nx << 16);
                lens_host[s] = n;
            }
        }
        __syncwarp();  // the staged abscissae are complete before the value line looks them up
    }
}

// writes the records at the word offsets off[] (prefix sums of wsize over the chunk's slots)
__global__ void __launch_bounds__(kCT) compact_write_kernel(DeviceBatch B, size_t first_pair, size_t npairs, const int32_t* __restrict__ aux,
                                                            const int32_t* __restrict__ yws, const unsigned* __restrict__ codes,
                                                            const long long* __restrict__ off, double* __restrict__ dst) {
    const int wid = threadIdx.x >> 5;
    const size_t lq = (size_t)blockIdx.x * kPairsPerCta + wid;
    if (lq >= npairs) return;
    const size_t q = first_pair + lq;
    const size_t sc = 2 * q, sv = sc + 1;
    const int cap = B.lay.cap;
    int nc = B.lens[sc], nv = B.lens[sv];
    nc = nc < cap ? nc : cap;
    nv = nv < cap ? nv : cap;
    if (yws[sc - 2 * first_pair] == kAliasMark) return;  // no record of its own (compact_count_kernel)
    ClassifyCtx cx{};  // the write pass classifies nothing
    const double2* Lc = B.lines + sc * (size_t)cap;
    const double2* Lv = B.lines + sv * (size_t)cap;
    for (int which = 0; which < 2; which++) {
        const int n = which ? nv : nc;
        if (n == 0) continue;
        const size_t s = (which ? sv : sc) - 2 * first_pair;
        const int ax = aux[s];
        LineSizes known;
        known.nrun = ax & 0xffff;
        known.nx = (ax >> 16) & 0xffff;
        known.yw = yws[s];
        encode_line<true>(which ? Lv : Lc, n, which == 1, cx, const_cast<unsigned*>(codes) + s * (size_t)cap,
                          known.nrun == 1 && known.nx == n, known, dst + off[s]);
    }
}

cudaError_t launch_compact_count(const DeviceBatch& B, int first, int count, bool alias, int32_t* aux, int32_t* yws, int32_t* wsize,
                                 unsigned* codes, int32_t* aux_host, int32_t* lens_host, cudaStream_t stream) {
    const size_t pairs_per = (size_t)B.lay.nT * B.lay.nstate * 2;
    if (count <= 0) return cudaSuccess;
    const size_t np = pairs_per * count;
    compact_count_kernel<<<(unsigned)((np + kPairsPerCta - 1) / kPairsPerCta), kCT, 0, stream>>>(B, pairs_per * first, np, alias ? pairs_per : 0,
                                                                                              aux, yws, wsize, codes, aux_host, lens_host);
    return cudaGetLastError();
}
cudaError_t launch_compact_write(const DeviceBatch& B, int first, int count, const int32_t* aux, const int32_t* yws, const unsigned* codes,
                                 const long long* off, double* dst, cudaStream_t stream) {
    const size_t pairs_per = (size_t)B.lay.nT * B.lay.nstate * 2;
    if (count <= 0) return cudaSuccess;
    const size_t np = pairs_per * count;
    compact_write_kernel<<<(unsigned)((np + kPairsPerCta - 1) / kPairsPerCta), kCT, 0, stream>>>(B, pairs_per * first, np, aux, yws, codes, off, dst);
    return cudaGetLastError();
}

}  // namespace dcegm
