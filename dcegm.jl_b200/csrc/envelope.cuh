// envelope.cuh — block-level secondary upper envelope + policy repair, all in shared memory.
//
// Replaces, for one (model, period[, income state]) worker step:
//   do_secondary        src/dc_algo.jl:393-495   (bend-back guard, kink insertion into the policy)
//   secondary_envelope  src/envelope.jl:268-285
//   splitLine           src/envelope.jl:80-126   (+ splitat, src/line.jl:323-345)
//   upper_env           src/envelope.jl:143-258  (+ interp extrap=false / linemax, line.jl:220-288)
//   removed!            src/envelope.jl:64-71
//   tail of the step    src/dc_algo.jl:94-101    (vbound, prepend, sortx!)
//
// The reference walks the union grid sequentially.  Here every grid point finds its sorted
// position, the winning section (first arg-max, strict >) and the interpolated top point with
// one binary search per section; changes of winner are handled independently (each one is a
// local root-finding problem on one cell of the union grid, with the reference's third-line
// loop), and the emission order of the reference is rebuilt with prefix sums.  The result is
// the same point sequence, intersection list and removed-index set as the sequential walk.
#pragma once
#include "device_common.cuh"

namespace dcegm {

// Block-strided loops of worker_envelope: kept rolled where the scratch is shared memory (the unrolled copies of ~40
// loops made the envelope 69 KB of code that every CTA streams through the instruction cache once per period),
// unrolled by four on long lines in global memory (independent loads in flight).
#ifndef DC_ENV_ROLLED
#define DC_ENV_ROLLED 1
#endif
#if DC_ENV_ROLLED
#define DC_ENV_UNROLL _Pragma("unroll (ES::kLongLines ? 4 : 1)")
#else
#define DC_ENV_UNROLL
#endif
// seg_interp_y: inline by default; solve_g_kernel (five envelopes per period, code footprint matters more than the
// call) compiles it out of line.  Measured: F 218.5 k vs 212.0 k models/s inline vs not, G 101.2 vs 98.1 ms.
#ifndef DC_SEG_INTERP_ATTR
#define DC_SEG_INTERP_ATTR __device__ __forceinline__
#endif

// All scratch of the envelope lives in the kernel's dynamic shared memory (dc_smem, declared in
// device_common.cuh) and is addressed as integer offsets from that one symbol.

template <class Mem>
struct EnvScratchT : Mem {
    int off;  // byte offset of the region inside the scratch memory (dc_smem, or the model's global region)
    int cap, isec_cap;
    // region layout: 5 double[cap] | 8 double[isec_cap] | redd[kWarps] | scan int[cap] | segb int[kSegCap+1]
    //                | 2 int[isec_cap] | red int[kWarps+1] | cnt int[8] | 5 u16[cap] | 4 u8[cap] | segrise u8[kSegCap]
    //                | 2 u8[isec_cap]
    __host__ __device__ static int align16(int b) { return (b + 15) / 16 * 16; }
    __host__ __device__ int segcap() const { return Mem::kSegCapFixed > 0 ? Mem::kSegCapFixed : cap; }
#ifndef DC_ENV_GUARDS
    __host__ __device__ int o_d(int j) const { return off + j * align16(cap * 8); }
    __host__ __device__ int o_i(int j) const { return o_d(5) + j * align16(isec_cap * 8); }
    __host__ __device__ int o_redd() const { return o_i(8); }
    __host__ __device__ int o_scan() const { return o_redd() + align16(Mem::kMaxWarps * 8); }
    __host__ __device__ int o_segb() const { return o_scan() + align16(cap * 4); }
    __host__ __device__ int o_key(int j) const { return o_segb() + align16(((Mem::kSegCapFixed > 0 ? Mem::kSegCapFixed : cap) + 1) * 4) + j * align16(isec_cap * 4); }
    __host__ __device__ int o_red() const { return o_key(2); }
    __host__ __device__ int o_cnt() const { return o_red() + align16((Mem::kMaxWarps + 1) * 4); }
    __host__ __device__ int o_u16(int j) const { return o_cnt() + 32 + j * align16(cap * 2); }
    __host__ __device__ int o_u8(int j) const { return o_u16(5) + j * align16(cap); }
    __host__ __device__ int o_segrise() const { return o_u8(4); }
    __host__ __device__ int o_ok(int j) const { return o_segrise() + align16(Mem::kSegCapFixed > 0 ? Mem::kSegCapFixed : cap) + j * align16(isec_cap); }
    __host__ __device__ int bytes() const { return o_ok(2) - off; }
#else
    // Debug build (-DDC_ENV_GUARDS, tools/env_guards_run.sh): 16 guard bytes behind every array.  worker_envelope fills
    // them (and the alignment padding) on entry and checks them on exit; a written guard shows up as status bit 15 — the
    // stand-in for compute-sanitizer's memcheck on the envelope's index arithmetic, in shared and in global memory
    static constexpr int kGuard = 16;
    __host__ __device__ static int slot(int b) { return align16(b) + kGuard; }
    __host__ __device__ int o_d(int j) const { return off + j * slot(cap * 8); }
    __host__ __device__ int o_i(int j) const { return o_d(5) + j * slot(isec_cap * 8); }
    __host__ __device__ int o_redd() const { return o_i(8); }
    __host__ __device__ int o_scan() const { return o_redd() + slot(Mem::kMaxWarps * 8); }
    __host__ __device__ int o_segb() const { return o_scan() + slot(cap * 4); }
    __host__ __device__ int o_key(int j) const { return o_segb() + slot((segcap() + 1) * 4) + j * slot(isec_cap * 4); }
    __host__ __device__ int o_red() const { return o_key(2); }
    __host__ __device__ int o_cnt() const { return o_red() + slot((Mem::kMaxWarps + 1) * 4); }
    __host__ __device__ int o_u16(int j) const { return o_cnt() + 32 + kGuard + j * slot(cap * 2); }
    __host__ __device__ int o_u8(int j) const { return o_u16(5) + j * slot(cap); }
    __host__ __device__ int o_segrise() const { return o_u8(4); }
    __host__ __device__ int o_ok(int j) const { return o_segrise() + slot(segcap()) + j * slot(isec_cap); }
    __host__ __device__ int bytes() const { return o_ok(2) - off; }
#endif
#ifdef DC_ENV_GUARDS
    // f(offset of an array, its bytes, bytes of its slot) for every array of the region
    template <class F>
    __device__ void each_array(F f) const {
        for (int j = 0; j < 5; j++) f(o_d(j), cap * 8, slot(cap * 8));
        for (int j = 0; j < 8; j++) f(o_i(j), isec_cap * 8, slot(isec_cap * 8));
        f(o_redd(), Mem::kMaxWarps * 8, slot(Mem::kMaxWarps * 8));
        f(o_scan(), cap * 4, slot(cap * 4));
        f(o_segb(), (segcap() + 1) * 4, slot((segcap() + 1) * 4));
        for (int j = 0; j < 2; j++) f(o_key(j), isec_cap * 4, slot(isec_cap * 4));
        f(o_red(), (Mem::kMaxWarps + 1) * 4, slot((Mem::kMaxWarps + 1) * 4));
        f(o_cnt(), 32, 32 + kGuard);
        for (int j = 0; j < 5; j++) f(o_u16(j), cap * 2, slot(cap * 2));
        for (int j = 0; j < 4; j++) f(o_u8(j), cap, slot(cap));
        f(o_segrise(), segcap(), slot(segcap()));
        for (int j = 0; j < 2; j++) f(o_ok(j), isec_cap, slot(isec_cap));
    }
    __device__ void guards_fill() const {
        unsigned char* base = reinterpret_cast<unsigned char*>(this->mem());
        each_array([&](int o, int b, int sl) {
            for (int k = b; k < sl; k++) base[o + k] = 0xA5;
        });
    }
    __device__ int guards_bad() const {  // arrays whose guard was written
        const unsigned char* base = reinterpret_cast<const unsigned char*>(this->mem());
        int bad = 0;
        each_array([&](int o, int b, int sl) {
            int hit = 0;
            for (int k = b; k < sl; k++) hit |= base[o + k] != 0xA5;
            bad += hit;
        });
        return bad;
    }
#endif
#define DC_ARR(T, name, expr) \
    __device__ __forceinline__ T* name() const { return reinterpret_cast<T*>(this->mem() + (expr)); }
    // extended dirty worker line (EGM order): x, value, policy.  The EGM step writes the na points at
    // offset `koff` = max(na/10-1,0); bend-back points are prepended in front.
    DC_ARR(double, Dx, o_d(0)) DC_ARR(double, Dv, o_d(1)) DC_ARR(double, Dc, o_d(2))
    // per point: interpolated top Point of the winning section at this x
    DC_ARR(double, Tx, o_d(3)) DC_ARR(double, Ty, o_d(4))
    // intersections (unsorted / sorted by emission order), policy points to insert (dc_algo.jl:434-479)
    DC_ARR(double, islx, o_i(0)) DC_ARR(double, isly, o_i(1)) DC_ARR(double, isl2x, o_i(2)) DC_ARR(double, isl2y, o_i(3))
    DC_ARR(double, insLx, o_i(4)) DC_ARR(double, insLy, o_i(5)) DC_ARR(double, insRx, o_i(6)) DC_ARR(double, insRy, o_i(7))
    DC_ARR(double, redd, o_redd())
    DC_ARR(int, scan, o_scan()) DC_ARR(int, segb, o_segb()) DC_ARR(int, islkey, o_key(0)) DC_ARR(int, isl2key, o_key(1))
    DC_ARR(int, red, o_red()) DC_ARR(int, cnt, o_cnt())  // cnt: [0] nsw/nl, [1] nis/nr, [2] status bits
    DC_ARR(unsigned short, pos, o_u16(0)) DC_ARR(unsigned short, order, o_u16(1)) DC_ARR(unsigned short, ord2, o_u16(2))
    DC_ARR(unsigned short, win, o_u16(3)) DC_ARR(unsigned short, swj, o_u16(4))
    DC_ARR(unsigned char, fknot, o_u8(0)) DC_ARR(unsigned char, fuq, o_u8(1)) DC_ARR(unsigned char, frem, o_u8(2))
    DC_ARR(unsigned char, cntI, o_u8(3)) DC_ARR(unsigned char, segrise, o_segrise())
    DC_ARR(unsigned char, insLok, o_ok(0)) DC_ARR(unsigned char, insRok, o_ok(1))
};
using EnvScratch = EnvScratchT<MemShared>;
using EnvScratchG = EnvScratchT<MemGlobal>;

struct SegView {
    const double* X;
    const double* Y;
    int base, st, len;
};
template <class ES>
__device__ __forceinline__ SegView seg_view(const ES& S, const double* X, const double* Y, int s) {
    SegView v;
    int b0 = S.segb()[s], b1 = S.segb()[s + 1];
    bool rise = S.segrise()[s];
    v.X = X;
    v.Y = Y;
    v.base = rise ? b0 : b1;
    v.st = rise ? 1 : -1;
    v.len = b1 - b0 + 1;
    return v;
}
__device__ __forceinline__ double seg_x(const SegView& v, int i) { return v.X[v.base + v.st * i]; }
__device__ __forceinline__ double seg_y(const SegView& v, int i) { return v.Y[v.base + v.st * i]; }
__device__ __forceinline__ bool seg_in_domain(const SegView& v, double q) {
    return !(q < seg_x(v, 0) || q > seg_x(v, v.len - 1));
}
// #knots <= q on the ascending view
__device__ __forceinline__ int seg_count_le(const SegView& v, double q) {
    int lo = 0, len = v.len;
    while (len > 0) {
        int half = len >> 1;
        bool le = seg_x(v, lo + half) <= q;
        lo = le ? lo + half + 1 : lo;
        len = le ? len - half - 1 : half;
    }
    return lo;
}
// interp(L[s],[q]) with extrap=true (line.jl:220-254), y only
DC_SEG_INTERP_ATTR double seg_interp_y(const SegView& v, double q) {
    const int n = v.len;
    if (q < seg_x(v, 0)) {
        double d = seg_x(v, 1) - seg_x(v, 0);
        return seg_y(v, 0) + ((seg_y(v, 1) - seg_y(v, 0)) / d) * (q - seg_x(v, 0));
    }
    if (q > seg_x(v, n - 1)) {
        double d = seg_x(v, n - 1) - seg_x(v, n - 2);
        return seg_y(v, n - 1) + ((seg_y(v, n - 1) - seg_y(v, n - 2)) / d) * (q - seg_x(v, n - 1));
    }
    int i = seg_count_le(v, q) - 1;
    i = i < 0 ? 0 : i;
    i = i > n - 2 ? n - 2 : i;
    double f = (q - seg_x(v, i)) / (seg_x(v, i + 1) - seg_x(v, i));
    return (1.0 - f) * seg_y(v, i) + f * seg_y(v, i + 1);
}

// One change of winner between union-grid cells j-1 and j (envelope.jl:174-243), resolved by one
// warp: all lanes follow the same control flow, the scan over all sections is split across lanes.
// Returns the number of intersections found; points go to px/py (at most kSwitchMaxIsect are stored).
template <class ES>
__device__ inline int resolve_switch(const ES& S, const double* X, const double* V, int nseg, int k0, int k1,
                                     double xl, double xr, double* px, double* py, unsigned* status) {
    const int lane = threadIdx.x & 31;
    int ln1 = k0, ln2 = k1;
    double pt1 = xl, pt2 = xr;
    SegView A = seg_view(S, X, V, ln1), B = seg_view(S, X, V, ln2);
    bool neither = seg_in_domain(A, pt1) && seg_in_domain(A, pt2) && seg_in_domain(B, pt1) && seg_in_domain(B, pt2);
    double fa = seg_interp_y(A, pt1) - seg_interp_y(B, pt1);
    double fb = seg_interp_y(A, pt2) - seg_interp_y(B, pt2);
    bool diffsign = fa * fb < 0;
    int cnt = 0;
    if (!(neither && diffsign)) return 0;
    int guard = 0;
    while (true) {
        if (++guard > 4 * nseg + 16) {
            *status |= DCEGM_ST_LOOP_LIMIT;
            break;
        }
        A = seg_view(S, X, V, ln1);
        B = seg_view(S, X, V, ln2);
        // root of A-B in [pt1,pt2]: both are linear inside one cell of the union grid (envelope.jl:199)
        double f1 = seg_interp_y(A, pt1) - seg_interp_y(B, pt1);
        double f2 = seg_interp_y(A, pt2) - seg_interp_y(B, pt2);
        double pt3;
        if (f1 == 0.0)
            pt3 = pt1;
        else if (f2 == 0.0)
            pt3 = pt2;
        else if (!(f1 * f2 < 0.0)) {
            *status |= DCEGM_ST_NO_BRACKET;
            break;
        } else {
            pt3 = pt1 - f1 * (pt2 - pt1) / (f2 - f1);
            if (pt3 < pt1) pt3 = pt1;
            if (pt3 > pt2) pt3 = pt2;
        }
        // all sections at pt3, extrap=false, first arg-max (envelope.jl:205-206): the lanes of the warp
        // share the sections; ties go to the lowest index and a NaN only wins as section 0, exactly
        // like the sequential scan `if (ln3 < 0 || y > y3)`
        int ln3 = 0x7fffffff;
        double y3 = -CUDART_INF, y_first = 0.0;
        for (int s = lane; s < nseg; s += 32) {
            SegView C = seg_view(S, X, V, s);
            double y = seg_in_domain(C, pt3) ? seg_interp_y(C, pt3) : -CUDART_INF;
            if (s == 0) y_first = y;
            if (y != y) y = -CUDART_INF;
            if (y > y3 || ln3 == 0x7fffffff) {
                y3 = y;
                ln3 = s;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double yo = __shfl_xor_sync(0xffffffffu, y3, o);
            const int so = __shfl_xor_sync(0xffffffffu, ln3, o);
            if (yo > y3 || (yo == y3 && so < ln3)) {
                y3 = yo;
                ln3 = so;
            }
        }
        y_first = __shfl_sync(0xffffffffu, y_first, 0);
        if (y_first != y_first) {
            ln3 = 0;
            y3 = y_first;
        }
        if (ln3 == ln2 || ln3 == ln1) {
            if (cnt < kSwitchMaxIsect) {
                px[cnt] = pt3;
                py[cnt] = y3;
            } else
                *status |= DCEGM_ST_ISECT_OVERFLOW;
            cnt++;
            if (ln2 == k1) break;
            ln1 = ln2;
            pt1 = pt3;
            ln2 = k1;
            pt2 = xr;
        } else {
            ln2 = ln3;
            pt2 = pt3;
        }
    }
    return cnt < kSwitchMaxIsect ? cnt : kSwitchMaxIsect;
}

// Full worker step after the EGM inversion.  On entry the na dirty points sit in
// S.Dx()/S.Dv()/S.Dc()[koff .. koff+na).  On exit (Vx,Vy,*nV) holds m.v[1,it].env and (Cx,Cy,*nC)
// m.c[1,it].env, both with the prepended credit-constraint point; meta / isects / removed are
// written to global memory.  All threads of the block must call this.
template <class ES, int NT = kThreads>
__device__ inline void worker_envelope(const ES& S, int koff, int na, double a0, double ev0, const DevModel& p,
                                       double* Vx, double* Vy, int* nV, double* Cx, double* Cy, int* nC,
                                       dcegm_env_meta* meta_out, double2* isects_out, uint32_t* removed_out,
                                       int rem_words, bool first_period = false) {
    const int tid = dc_gtid<NT>();  // index inside the thread group (the CTA unless DC_GROUPED)
    unsigned status = 0;
#ifdef DC_ENV_GUARDS
    if (tid == 0) S.guards_fill();  // ordered before every write of this call by the barrier below
#ifdef DC_ENV_GUARD_SELFTEST  // positive control: one int written one element behind an array must be reported
    if (tid == 1) S.scan()[S.cap] = 1;
#endif
#endif

    // ---- bend-back guard (dc_algo.jl:394-418) ------------------------------------------------
    double mn = CUDART_INF;
    DC_ENV_UNROLL
    for (int k = tid; k < na; k += NT) mn = fmin(mn, S.Dx()[koff + k]);
    const double minx = block_min<NT>(mn, S.redd());
    const double first = S.Dx()[koff];
    int npre = 0;
    if (!(first <= minx)) {
        npre = koff;  // floor(na/10) - 1 points
        const int klen = na / 10;
        DC_ENV_UNROLL
        for (int i = tid; i < npre; i += NT) {
            double x0 = dc_linrange_at(minx, first, klen, i);
            double c0 = x0;
            if (minx < 0) c0 = c0 + fabs(minx) + p.cfloor;
            S.Dx()[i] = x0;
            S.Dc()[i] = c0;
            S.Dv()[i] = u_work(c0, true, p) + p.beta * ev0;
        }
    }
    const int s0 = koff - npre;
    const double* X = S.Dx() + s0;
    const double* V = S.Dv() + s0;
    const double* Cc = S.Dc() + s0;
    const int n = na + npre;
    DC_CLK(4);
    if (tid == 0) {
        S.cnt()[0] = 0;
        S.cnt()[1] = 0;
        S.cnt()[2] = 0;
    }
    dc_gsync<NT>();

    // ---- splitLine: direction changes (envelope.jl:84-110) -----------------------------------
    DC_ENV_UNROLL
    for (int k = tid; k < n; k += NT) {
        int isb = 0;
        if (k >= 1 && k <= n - 2) isb = (X[k] > X[k - 1]) != (X[k + 1] > X[k]);
        S.scan()[k] = isb;
        S.fknot()[k] = (unsigned char)isb;
    }
    dc_gsync<NT>();
    const int nb = block_scan_excl<NT>(S.scan(), n, S.red());
    int nseg = nb + 1;
    const bool first_rising = X[1] > X[0];
    bool plain = false;
    if (nseg == 1) {
        plain = true;
        if (!first_rising) {
            status |= DCEGM_ST_ALL_FALLING;
            nseg = 0;
        }
    } else if (nseg > (ES::kSegCapFixed > 0 ? ES::kSegCapFixed : S.cap)) {
        status |= DCEGM_ST_SEG_OVERFLOW;
        plain = true;
    }

    int ne = 0, nis = 0, nrem = 0;
    if (plain) {
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) {
            Vx[1 + k] = X[k];
            Vy[1 + k] = V[k];
            S.frem()[k] = 0;
        }
        ne = n;
    } else {
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT)
            if (S.fknot()[k]) S.segb()[S.scan()[k] + 1] = k;
        if (tid == 0) {
            S.segb()[0] = 0;
            S.segb()[nseg] = n - 1;
        }
        dc_gsync<NT>();
        DC_ENV_UNROLL
        for (int s = tid; s < nseg; s += NT) S.segrise()[s] = X[S.segb()[s] + 1] > X[S.segb()[s]];
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) S.order()[k] = 0;
        dc_gsync<NT>();

        DC_CLK(5);
        // ---- every point: sorted position, winning section, top Point (envelope.jl:154-164) ----
        // fine-grid solver: the extent and the size of the first kSegTile sections are staged in shared memory, so
        // that the test below costs two shared loads instead of four dependent global ones
        constexpr bool kStage = ES::kLongLines;
        constexpr int kSegTile = kStage ? 1536 : 1;
        __shared__ double t_lo[kSegTile], t_hi[kSegTile];
        __shared__ int t_own[kSegTile];
        // ... and per group of 32 consecutive sections their common extent and total size: neighbouring sections
        // are neighbours in x (small wiggles of the value line), so most groups are skipped whole
        constexpr int kGroups = kSegTile / 32 + 1;
        __shared__ double g_lo[kGroups], g_hi[kGroups];
        __shared__ int g_own[kGroups];
        const int ntile = kStage ? (nseg < kSegTile ? nseg : kSegTile) : 0;
        if (kStage) {
            DC_ENV_UNROLL
            for (int s = tid; s < ntile; s += NT) {
                const int e0 = S.segb()[s], e1 = S.segb()[s + 1];
                const double xa = X[e0], xb = X[e1];
                t_lo[s] = xa < xb ? xa : xb;
                t_hi[s] = xa > xb ? xa : xb;
                t_own[s] = e1 - e0 + (s == 0 ? 1 : 0);
            }
            dc_gsync<NT>();
            for (int gi = tid; gi * 32 + 32 <= ntile; gi += NT) {
                double lo = CUDART_INF, hi = -CUDART_INF;
                int own = 0;
                for (int s = gi * 32; s < gi * 32 + 32; s++) {
                    const double a = t_lo[s], b = t_hi[s];
                    lo = (a == a) ? (a < lo ? a : lo) : -CUDART_INF;  // a NaN extent is never skipped
                    hi = (b == b) ? (b > hi ? b : hi) : CUDART_INF;
                    own += t_own[s];
                }
                g_lo[gi] = lo;
                g_hi[gi] = hi;
                g_own[gi] = own;
            }
            dc_gsync<NT>();
        }
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) {
            const double xk = X[k];
            int rank = 0, ties = 0, w = -1;
            double best = -CUDART_INF, bestx = xk;
            bool knot = false;
            for (int s = 0; s < nseg; s++) {
                // sections entirely to one side of xk need no search: left of xk they contribute all their
                // own points to the rank (the boundary shared with the previous section is not theirs),
                // right of xk nothing; neither contains xk
                if (kStage && s < ntile) {
                    if ((s & 31) == 0 && s + 32 <= ntile) {  // a whole group on one side of xk
                        if (xk > g_hi[s >> 5]) {
                            rank += g_own[s >> 5];
                            s += 31;
                            continue;
                        }
                        if (xk < g_lo[s >> 5]) {
                            s += 31;
                            continue;
                        }
                    }
                    if (xk > t_hi[s]) {
                        rank += t_own[s];
                        continue;
                    }
                    if (xk < t_lo[s]) continue;
                } else {
                    const int e0 = S.segb()[s], e1 = S.segb()[s + 1];
                    const double xa = X[e0], xb = X[e1];
                    if (xk > (xa > xb ? xa : xb)) {
                        rank += e1 - e0 + (s == 0 ? 1 : 0);
                        continue;
                    }
                    if (xk < (xa < xb ? xa : xb)) continue;
                }
                SegView v = seg_view(S, X, V, s);
                // #knots < xk
                int lo = 0, len = v.len;
                while (len > 0) {
                    int half = len >> 1;
                    bool lt = seg_x(v, lo + half) < xk;
                    lo = lt ? lo + half + 1 : lo;
                    len = lt ? len - half - 1 : half;
                }
                const int cnt = lo;
                const bool eq = cnt < v.len && seg_x(v, cnt) == xk;
                // each point is owned by one section: the shared boundary belongs to the earlier one
                const int b0 = S.segb()[s];
                int own = cnt;
                if (s > 0 && X[b0] < xk) own -= 1;
                rank += own;
                if (eq) {
                    int i = v.base + v.st * cnt;
                    if ((s == 0 || i != b0) && i < k) ties++;
                }
                if (!(xk < seg_x(v, 0) || xk > seg_x(v, v.len - 1))) {
                    int i = cnt + (eq ? 1 : 0) - 1;
                    i = i < 0 ? 0 : i;
                    i = i > v.len - 2 ? v.len - 2 : i;
                    const double xa = seg_x(v, i), xb = seg_x(v, i + 1);
                    const double f = (xk - xa) / (xb - xa);
                    const double yv = (1.0 - f) * seg_y(v, i) + f * seg_y(v, i + 1);
                    if (yv > best) {  // linemax: first strictly larger wins (line.jl:280)
                        best = yv;
                        bestx = (1.0 - f) * xa + f * xb;
                        w = s;
                        knot = (xa == xk) || (xb == xk);
                    }
                }
            }
            if (w < 0) {
                status |= DCEGM_ST_NONFINITE;
                w = 0;
            }
            int ps = rank + ties;
            ps = ps < n ? ps : n - 1;
            S.pos()[k] = (unsigned short)ps;
            S.win()[k] = (unsigned short)w;
            S.Tx()[k] = bestx;
            S.Ty()[k] = best;
            S.fknot()[k] = knot;
        }
        dc_gsync<NT>();
        DC_CLK(6);
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) S.order()[S.pos()[k]] = (unsigned short)k;
        dc_gsync<NT>();
        // unique(xx)
        DC_ENV_UNROLL
        for (int q = tid; q < n; q += NT) {
            int uq = (q == 0) || (X[S.order()[q]] != X[S.order()[q - 1]]);
            S.fuq()[q] = (unsigned char)uq;
            S.scan()[q] = uq;
        }
        dc_gsync<NT>();
        const int nu = block_scan_excl<NT>(S.scan(), n, S.red());
        DC_ENV_UNROLL
        for (int q = tid; q < n; q += NT)
            if (S.fuq()[q]) S.ord2()[S.scan()[q]] = S.order()[q];
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) {
            int q = S.pos()[k];
            S.pos()[k] = (unsigned short)(S.fuq()[q] ? S.scan()[q] : S.scan()[q] - 1);  // index into the unique grid
        }
        dc_gsync<NT>();

        DC_CLK(7);
        // ---- changes of winner (envelope.jl:173-243) ---------------------------------------------
        DC_ENV_UNROLL
        for (int j = tid; j < nu; j += NT) {
            S.cntI()[j] = 0;
            if (j >= 1 && S.win()[S.ord2()[j]] != S.win()[S.ord2()[j - 1]]) {
                int slot = atomicAdd(&S.cnt()[0], 1);
                S.swj()[slot] = (unsigned short)j;
            }
        }
        dc_gsync<NT>();
        const int nsw = S.cnt()[0];
        for (int t = tid >> 5; t < nsw; t += (NT / 32)) {  // warp per change of winner
            const int j = S.swj()[t];
            double px[kSwitchMaxIsect], py[kSwitchMaxIsect];
            int c = resolve_switch(S, X, V, nseg, S.win()[S.ord2()[j - 1]], S.win()[S.ord2()[j]], X[S.ord2()[j - 1]], X[S.ord2()[j]],
                                   px, py, &status);
            if ((tid & 31) != 0) c = 0;  // lane 0 publishes
            if (c > 0) {
                int base = atomicAdd(&S.cnt()[1], c);
                for (int q = 0; q < c; q++) {
                    if (base + q < S.isec_cap) {
                        S.islx()[base + q] = px[q];
                        S.isly()[base + q] = py[q];
                        S.islkey()[base + q] = j * kSwitchMaxIsect + q;
                    }
                }
                S.cntI()[j] = (unsigned char)c;
            }
        }
        dc_gsync<NT>();
        DC_CLK(8);
        nis = S.cnt()[1];
        if (nis > S.isec_cap) {
            status |= DCEGM_ST_ISECT_OVERFLOW;
            nis = S.isec_cap;
        }
        // intersections in emission order (ascending cell, then order of discovery)
        DC_ENV_UNROLL
        for (int t = tid; t < nis; t += NT) {
            int key = S.islkey()[t], r = 0;
            for (int u = 0; u < nis; u++) r += S.islkey()[u] < key;
            S.isl2x()[r] = S.islx()[t];
            S.isl2y()[r] = S.isly()[t];
            S.isl2key()[r] = key;
        }
        // ---- emission (envelope.jl:170, 212, 248-252) ----------------------------------------------
        DC_ENV_UNROLL
        for (int j = tid; j < nu; j += NT) {
            int k = S.ord2()[j];
            int reg = (j == 0) || S.fknot()[k] || (j == nu - 1);
            S.scan()[j] = (j == 0 ? 0 : (int)S.cntI()[j]) + reg;
        }
        dc_gsync<NT>();
        ne = block_scan_excl<NT>(S.scan(), nu, S.red());
        const int room = S.cap - 1;
        if (ne > room) status |= DCEGM_ST_LINE_OVERFLOW;
        DC_ENV_UNROLL
        for (int j = tid; j < nu; j += NT) {
            int k = S.ord2()[j];
            int reg = (j == 0) || S.fknot()[k] || (j == nu - 1);
            int at = S.scan()[j] + (j == 0 ? 0 : (int)S.cntI()[j]);
            if (reg && at < room) {
                Vx[1 + at] = S.Tx()[k];
                Vy[1 + at] = S.Ty()[k];
            }
        }
        DC_ENV_UNROLL
        for (int t = tid; t < nis; t += NT) {
            int key = S.isl2key()[t];
            int at = S.scan()[key / kSwitchMaxIsect] + key % kSwitchMaxIsect;
            if (at < room) {
                Vx[1 + at] = S.isl2x()[t];
                Vy[1 + at] = S.isl2y()[t];
            }
        }
        if (ne > room) ne = room;
        // ---- removed! (envelope.jl:64-71): points with no exactly equal Point in the envelope ------
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) {
            int j = S.pos()[k];
            bool reg = (j == 0) || S.fknot()[k] || (j == nu - 1);
            bool kept = reg && S.Tx()[k] == X[k] && S.Ty()[k] == V[k];
            for (int t = 0; t < nis && !kept; t++) kept = (S.isl2x()[t] == X[k]) && (S.isl2y()[t] == V[k]);
            S.frem()[k] = kept ? 0 : 1;
        }
        dc_gsync<NT>();
    }
    DC_CLK(9);
    {
        int local = 0;
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) local += S.frem()[k];
        nrem = block_sum<NT>(local, S.red());
    }
    // removed bitmask -> global
    for (int base = 0; base < rem_words * 32; base += NT) {
        int k = base + tid;
        unsigned bit = (k < n) ? S.frem()[k] : 0;
        unsigned word = __ballot_sync(0xffffffffu, bit);
        if ((tid & 31) == 0 && k / 32 < rem_words) removed_out[k / 32] = word;
    }
    DC_ENV_UNROLL
    for (int t = tid; t < S.isec_cap; t += NT)
        isects_out[t] = t < nis ? make_double2(S.isl2x()[t], S.isl2y()[t]) : make_double2(0.0, 0.0);

    DC_CLK(5);
    // ---- policy repair (dc_algo.jl:423-493) -----------------------------------------------------
    int ncur;
    if (nrem > 0) {
        const int lane = tid & 31, wid = tid >> 5;
        for (int t = wid; t < nis; t += (NT / 32)) {
            const double ix = S.isl2x()[t];
            int jl = -1, jr = 0x7fffffff;
            for (int k = lane; k < n; k += 32) {
                if (!S.frem()[k]) {
                    if (X[k] < ix) jl = max(jl, k);
                    if (X[k] > ix) jr = min(jr, k);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                jl = max(jl, __shfl_xor_sync(0xffffffffu, jl, o));
                jr = min(jr, __shfl_xor_sync(0xffffffffu, jr, o));
            }
            if (lane == 0) {
                S.insLok()[t] = 0;
                S.insRok()[t] = 0;
                if (jl >= 0 && jl < n - 1) {  // dc_algo.jl:446-460
                    double2 r = interp2pt(make_double2(X[jl], Cc[jl]), make_double2(X[jl + 1], Cc[jl + 1]), ix);
                    S.insLx()[t] = r.x - 1e3 * 2.220446049250313e-16;
                    S.insLy()[t] = r.y;
                    S.insLok()[t] = 1;
                }
                if (jr != 0x7fffffff) {  // dc_algo.jl:468-476
                    if (jr == 0) {
                        atomicOr(&S.cnt()[2], (int)DCEGM_ST_INSERT_RANGE);
                    } else {
                        double2 r = interp2pt(make_double2(X[jr - 1], Cc[jr - 1]), make_double2(X[jr], Cc[jr]), ix);
                        S.insRx()[t] = r.x;
                        S.insRy()[t] = r.y;
                        S.insRok()[t] = 1;
                    }
                }
            }
        }
        // deleteat!(cout, rmidx)  (dc_algo.jl:482)
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) S.scan()[k] = S.frem()[k] ? 0 : 1;
        dc_gsync<NT>();
        // compact the two insertion lists (push! order)
        if (tid == 0) {
            int nl = 0, nr = 0;
            for (int t = 0; t < nis; t++) {
                if (S.insLok()[t]) {
                    S.insLx()[nl] = S.insLx()[t];
                    S.insLy()[nl] = S.insLy()[t];
                    nl++;
                }
                if (S.insRok()[t]) {
                    S.insRx()[nr] = S.insRx()[t];
                    S.insRy()[nr] = S.insRy()[t];
                    nr++;
                }
            }
            S.cnt()[0] = nl;
            S.cnt()[1] = nr;
        }
        ncur = block_scan_excl<NT>(S.scan(), n, S.red());
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT)
            if (!S.frem()[k]) {
                Cx[1 + S.scan()[k]] = X[k];
                Cy[1 + S.scan()[k]] = Cc[k];
            }
        dc_gsync<NT>();
        const int nl = S.cnt()[0], nr = S.cnt()[1];
        // the reference's loop runs over insert_left and indexes insert_right with it (dc_algo.jl:486-492): a shorter
        // insert_right is a BoundsError, a longer one is silently ignored
        if (nr < nl) status |= DCEGM_ST_MISALIGNED;
        if (nr > nl) status |= DCEGM_ST_EXTRA_RIGHT;
        // insert!(…) pairs (dc_algo.jl:486-492).  The reference inserts one pair after the other, each before
        // the first x greater than its left point in the line as it stands.  When the surviving policy is
        // ascending and every pair lies at or right of the pair before it (the regular case: intersections come
        // in ascending order), that sequence equals one merge: pair i lands in front of base point
        // q_i = #{x <= left_i}, behind the i pairs before it.  Anything else takes the sequential path below.
        bool merged = false;
        constexpr bool kTryMerge = ES::kLongLines;  // at ~500 points the sequential path is as fast (measured)
        if (kTryMerge && nl > 0 && nl <= nr && ncur + 2 * nl <= S.cap - 1) {
            int irregular = 0;
            DC_ENV_UNROLL
            for (int k = tid; k + 1 < ncur; k += NT) irregular |= Cx[1 + k + 1] < Cx[1 + k];
            DC_ENV_UNROLL
            for (int i = tid; i < nl; i += NT) {
                const double lx = S.insLx()[i];
                int lo = 0, len = ncur;  // #{base x <= lx}
                while (len > 0) {
                    const int half = len >> 1;
                    const bool le = Cx[1 + lo + half] <= lx;
                    lo = le ? lo + half + 1 : lo;
                    len = le ? len - half - 1 : half;
                }
                S.islkey()[i] = lo;
                irregular |= lo >= ncur;  // no x greater than the left point: the reference raises here
                irregular |= !(lx == lx);
                if (i > 0) {
                    const double pl = S.insLx()[i - 1], pr = S.insRx()[i - 1];
                    irregular |= !(pl <= lx && pr <= lx);
                }
            }
            merged = block_sum<NT>(irregular, S.red()) == 0;
        }
        if (merged) {
            DC_ENV_UNROLL
            for (int k = tid; k < ncur; k += NT) {
                int lo = 0, len = nl;  // pairs in front of base point k: #{q_i <= k}
                while (len > 0) {
                    const int half = len >> 1;
                    const bool le = S.islkey()[lo + half] <= k;
                    lo = le ? lo + half + 1 : lo;
                    len = le ? len - half - 1 : half;
                }
                S.Dx()[k + 2 * lo] = Cx[1 + k];
                S.Dv()[k + 2 * lo] = Cy[1 + k];
            }
            DC_ENV_UNROLL
            for (int i = tid; i < nl; i += NT) {
                const int at = S.islkey()[i] + 2 * i;
                S.Dx()[at] = S.insLx()[i];
                S.Dv()[at] = S.insLy()[i];
                S.Dx()[at + 1] = S.insRx()[i];
                S.Dv()[at + 1] = S.insRy()[i];
            }
            dc_gsync<NT>();
            ncur += 2 * nl;
            DC_ENV_UNROLL
            for (int k = tid; k < ncur; k += NT) {
                Cx[1 + k] = S.Dx()[k];
                Cy[1 + k] = S.Dv()[k];
            }
            dc_gsync<NT>();
        }
        for (int i = 0; i < nl && !merged; i++) {
            const double lx = S.insLx()[i];
            int cand = 0x7fffffff;
            DC_ENV_UNROLL
            for (int k = tid; k < ncur; k += NT)
                if (Cx[1 + k] > lx) {
                    cand = k;
                    break;
                }
            const int j = block_min_int<NT>(cand, S.red());
            if (j == 0x7fffffff || i >= nr) {
                status |= DCEGM_ST_INSERT_RANGE;
                break;
            }
            if (ncur + 2 > S.cap - 1) {
                status |= DCEGM_ST_LINE_OVERFLOW;
                break;
            }
            // shift [j, ncur) right by two, highest chunk first
            for (int top = ncur - 1; top >= j; top -= NT) {
                int k = top - tid;
                double vx = 0, vy = 0;
                if (k >= j) {
                    vx = Cx[1 + k];
                    vy = Cy[1 + k];
                }
                dc_gsync<NT>();
                if (k >= j) {
                    Cx[1 + k + 2] = vx;
                    Cy[1 + k + 2] = vy;
                }
                dc_gsync<NT>();
            }
            if (tid == 0) {
                Cx[1 + j] = lx;
                Cy[1 + j] = S.insLy()[i];
                Cx[1 + j + 1] = S.insRx()[i];
                Cy[1 + j + 1] = S.insRy()[i];
            }
            ncur += 2;
            dc_gsync<NT>();
        }
    } else {
        DC_ENV_UNROLL
        for (int k = tid; k < n; k += NT) {
            Cx[1 + k] = X[k];
            Cy[1 + k] = Cc[k];
        }
        ncur = n;
    }

    DC_CLK(10);
    // ---- tail of the step (dc_algo.jl:94-101) ----------------------------------------------------
    if (tid == 0) {
        Vx[0] = a0;
        Vy[0] = ev0;
        Cx[0] = a0;
        Cy[0] = 0.0;
    }
    dc_gsync<NT>();
    const int lenV = ne + 1, lenC = ncur + 1;
    // sortx!(c): stable sort on x.  The repaired policy is already ascending except in degenerate
    // cases; detect, and only then rank-sort through the (now free) D arrays.
    int bad = 0;
    DC_ENV_UNROLL
    for (int k = tid; k + 1 < lenC; k += NT) bad |= Cx[k + 1] < Cx[k];
    bad = block_sum<NT>(bad, S.red());
    if (bad && ES::kLongLines) {
        // long lines (fine-grid solver): stable merge sort of the natural ascending runs, pairwise and
        // bottom-up; every element finds its place in the merged pair with one binary search in the sibling
        // run (left-run elements go before equal right-run elements).  Ping-pong with the free D arrays;
        // run boundaries in segb (room for cap + 1 entries in this variant), scan as temporary.
        DC_ENV_UNROLL
        for (int k = tid; k < lenC; k += NT) S.scan()[k] = (k == 0) || (Cx[k] < Cx[k - 1]);
        dc_gsync<NT>();
        DC_ENV_UNROLL
        for (int k = tid; k < lenC; k += NT) S.fuq()[k] = (unsigned char)S.scan()[k];
        dc_gsync<NT>();
        int nr = block_scan_excl<NT>(S.scan(), lenC, S.red());
        int* rb = S.segb();
        DC_ENV_UNROLL
        for (int k = tid; k < lenC; k += NT)
            if (S.fuq()[k]) rb[S.scan()[k]] = k;
        if (tid == 0) rb[nr] = lenC;
        dc_gsync<NT>();
        double *sx = Cx, *sy = Cy, *dx = S.Dx(), *dy = S.Dv();
        while (nr > 1) {
            DC_ENV_UNROLL
            for (int k = tid; k < lenC; k += NT) {
                int lo = 0, len = nr;  // run containing k: last boundary <= k
                while (len > 0) {
                    const int half = len >> 1;
                    const bool le = rb[lo + half] <= k;
                    lo = le ? lo + half + 1 : lo;
                    len = le ? len - half - 1 : half;
                }
                const int j = lo - 1;
                const double xk = sx[k];
                int out = k;
                if ((j & 1) == 0) {
                    if (j + 1 < nr) {  // count of x < xk in the right sibling
                        const int b0 = rb[j + 1];
                        int l2 = 0, n2 = rb[j + 2] - b0;
                        while (n2 > 0) {
                            const int half = n2 >> 1;
                            const bool lt = sx[b0 + l2 + half] < xk;
                            l2 = lt ? l2 + half + 1 : l2;
                            n2 = lt ? n2 - half - 1 : half;
                        }
                        out = k + l2;
                    }
                } else {  // count of x <= xk in the left sibling
                    const int b0 = rb[j - 1];
                    int l2 = 0, n2 = rb[j] - b0;
                    while (n2 > 0) {
                        const int half = n2 >> 1;
                        const bool le = sx[b0 + l2 + half] <= xk;
                        l2 = le ? l2 + half + 1 : l2;
                        n2 = le ? n2 - half - 1 : half;
                    }
                    out = b0 + l2 + (k - rb[j]);
                }
                dx[out] = xk;
                dy[out] = sy[k];
            }
            const int nr2 = (nr + 1) >> 1;
            DC_ENV_UNROLL
            for (int i = tid; i < nr2; i += NT) S.scan()[i] = rb[2 * i];
            dc_gsync<NT>();
            DC_ENV_UNROLL
            for (int i = tid; i < nr2; i += NT) rb[i] = S.scan()[i];
            if (tid == 0) rb[nr2] = lenC;
            dc_gsync<NT>();
            nr = nr2;
            double* t = sx; sx = dx; dx = t;
            t = sy; sy = dy; dy = t;
        }
        if (sx != Cx) {
            DC_ENV_UNROLL
            for (int k = tid; k < lenC; k += NT) {
                Cx[k] = sx[k];
                Cy[k] = sy[k];
            }
            dc_gsync<NT>();
        }
    } else if (bad) {
        DC_ENV_UNROLL
        for (int k = tid; k < lenC; k += NT) {
            const double xk = Cx[k];
            int r = 0;
            for (int i = 0; i < lenC; i++) {
                double xi = Cx[i];
                r += (xi < xk) || (xi == xk && i < k);
            }
            S.Dx()[r] = xk;
            S.Dv()[r] = Cy[k];
        }
        dc_gsync<NT>();
        DC_ENV_UNROLL
        for (int k = tid; k < lenC; k += NT) {
            Cx[k] = S.Dx()[k];
            Cy[k] = S.Dv()[k];
        }
        dc_gsync<NT>();
    }
    // status: knot vectors the previous period's interpolate() would reject (knot_pair_flag), non-finite entries
    int flag = 0;
    DC_ENV_UNROLL
    for (int k = tid; k < lenC; k += NT) {
        if (k + 1 < lenC) flag |= knot_pair_flag(Cx[k], Cx[k + 1], first_period);
        if (!isfinite(Cx[k]) || !isfinite(Cy[k])) flag |= DCEGM_ST_NONFINITE;
    }
    DC_ENV_UNROLL
    for (int k = tid; k < lenV; k += NT) {
        if (k + 1 < lenV) flag |= knot_pair_flag(Vx[k], Vx[k + 1], first_period);
        if (!isfinite(Vx[k]) || !isfinite(Vy[k])) flag |= DCEGM_ST_NONFINITE;
    }
    if (flag || status) atomicOr(&S.cnt()[2], flag | (int)status);
    dc_gsync<NT>();
    if (tid == 0) {
#ifdef DC_ENV_GUARDS
        if (S.guards_bad()) atomicOr(&S.cnt()[2], 1 << 15);  // an index left its array: shows up as status bit 15
#endif
        *nV = lenV;
        *nC = lenC;
        meta_out->n_segments = nseg;
        meta_out->n_isects = nis;
        meta_out->n_removed = nrem;
        meta_out->n_prepended = npre;
    }
    dc_gsync<NT>();
}

}  // namespace dcegm
