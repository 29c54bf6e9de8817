// compact_host.cpp — host-side decoder of the compact line format (compact_lines.cu, include/dcegm.h): rebuilds
// the (x, y) points of stored polylines — Vector{Point} of line.jl:1-4 — from ordinates + abscissa runs.
// Pure host code (no CUDA): one IEEE addition per GRID point, copies otherwise; the result is bit-identical to
// what dcegm_fetch_packed delivers (tests/test_gpu_compact.py).
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "host_setup.hpp"

namespace dcegm {

namespace {

inline uint64_t word_bits(double w) {
    uint64_t b;
    memcpy(&b, &w, 8);
    return b;
}

// The ordinate section (compact_lines.cu): I_0 | I_1 - I_0 | control bytes | payload; blocks of 8 points hold second or
// third differences of the 64-bit patterns, w four-bit units each.  Writes the n ordinates to out[2 k + 1].  false when
// the section is inconsistent.
template <int HB, bool ODD>
inline void decode_block(const unsigned char* src, const unsigned char* top, int j0, int cnt, bool third, uint64_t& i1,
                         uint64_t& d1, uint64_t& d2, double* dst) {
    constexpr int bits = 8 * HB + (ODD ? 4 : 0);  // width of a residual, 0 .. 64
    for (int j = j0; j < cnt; j++, dst += 2) {
        uint64_t r = 0;
        if constexpr (bits > 0) {
            uint64_t v = 0;
            // one unaligned 8-byte load per residual: the record's run words follow the payload, so the load never
            // leaves the record; bytes beyond the residual's own are shifted out below
            if constexpr (HB > 0) memcpy(&v, src + (size_t)j * HB, 8);
            if constexpr (ODD) v = (v & ((1ULL << (8 * HB)) - 1)) | ((uint64_t)((top[j >> 1] >> (4 * (j & 1))) & 15) << (8 * HB));
            r = (uint64_t)((int64_t)(v << (64 - bits)) >> (64 - bits));  // sign extension
        }
        d2 = third ? d2 + r : r;
        d1 += d2;
        i1 += d1;
        memcpy(dst, &i1, 8);
    }
}

bool decode_y(const double* sec, int yw, int n, double* out) {
    if (n <= 0) return yw == 0;
    const int nblk = (n + 7) / 8;  // block b: k = 8 b .. 8 b + 7; the residuals of k = 0, 1 are zero place-holders
    const int head = (n < 2 ? n : 2) + (nblk + 7) / 8;
    if (yw < head) return false;
    uint64_t i1 = word_bits(sec[0]), d1 = 0, d2 = 0;  // I_(k-1), I_(k-1) - I_(k-2), R2_(k-1)
    memcpy(&out[1], &i1, 8);
    if (n > 1) {
        d1 = word_bits(sec[1]);
        i1 += d1;
        memcpy(&out[3], &i1, 8);
    }
    const unsigned char* ctl = reinterpret_cast<const unsigned char*>(sec + (n < 2 ? n : 2));
    const unsigned char* pay = reinterpret_cast<const unsigned char*>(sec + head);
    const size_t pay_cap = (size_t)(yw - head) * 8;
    size_t at = 0;
    for (int blk = 0; blk < nblk; blk++) {
        const int c = ctl[blk], w = c & 31, hb = w >> 1;
        const bool third = (c & 32) != 0;
        const int cnt = n - blk * 8 < 8 ? n - blk * 8 : 8;
        const size_t bytes = (size_t)cnt * hb + ((w & 1) ? (cnt + 1) / 2 : 0);
        if (w > 16 || (c & 0xc0) || at + bytes > pay_cap) return false;
        const int k0 = blk * 8;
        const int j0 = k0 < 2 ? 2 - k0 : 0;  // the residuals of k = 0, 1 are place-holders (raw words in front)
        const unsigned char* src = pay + at;
        const unsigned char* top = src + (size_t)cnt * hb;
        double* dst = out + 2 * (size_t)(k0 + j0) + 1;
        switch (w) {
#define DC_W(W)                                                                                   \
    case W:                                                                                       \
        decode_block<(W) / 2, ((W) & 1) != 0>(src, top, j0, cnt, third, i1, d1, d2, dst);         \
        break;
            DC_W(0) DC_W(1) DC_W(2) DC_W(3) DC_W(4) DC_W(5) DC_W(6) DC_W(7) DC_W(8) DC_W(9) DC_W(10) DC_W(11) DC_W(12)
            DC_W(13) DC_W(14) DC_W(15) DC_W(16)
#undef DC_W
        }
        at += bytes;
    }
    return (at + 7) / 8 == (size_t)(yw - head);
}

// One slot: rec = head | ysec[yw] | run[nrun] | xs[nx].  xc/nc: decoded points (interleaved x, y) of the slot's policy
// line for REF runs (null for a policy line).  Writes n interleaved points to out.  false when the record is
// inconsistent.
bool decode_slot(const double* rec, int n, int nrun, int nx, const double* a, int na, const double* xc, int nc, double* out) {
    if (n == 0) return nrun == 0 && nx == 0;
    if (nrun < 1) return false;
    const uint64_t yw64 = word_bits(rec[0]);
    if (yw64 > (uint64_t)n + (uint64_t)n / 64 + 4) return false;
    const int yw = (int)yw64;
    if (!decode_y(rec + 1, yw, n, out)) return false;
    const double* runs = rec + 1 + yw;
    const double* xs = runs + nrun;
    int ix = 0;
    for (int r = 0; r < nrun; r++) {
        const uint64_t e = word_bits(runs[r]);
        const int pos = (int)(e & 0xffff), src = (int)((e >> 16) & 0xffff), kind = (int)((e >> 32) & 0xffff);
        const int end = r + 1 < nrun ? (int)(word_bits(runs[r + 1]) & 0xffff) : n;
        if ((r == 0 && pos != 0) || end <= pos || end > n) return false;
        const int len = end - pos;
        switch (kind) {
            case 0:  // GRID: x = a[i] + y
                if (src + len > na) return false;
                for (int j = 0; j < len; j++) out[2 * (pos + j)] = a[src + j] + out[2 * (pos + j) + 1];
                break;
            case 1:  // REF: the policy line's abscissae
                if (!xc || src + len > nc) return false;
                for (int j = 0; j < len; j++) out[2 * (pos + j)] = xc[2 * (src + j)];
                break;
            case 2:  // EXPLICIT
                if (ix + len > nx) return false;
                for (int j = 0; j < len; j++) out[2 * (pos + j)] = xs[ix++];
                break;
            default:
                return false;
        }
    }
    return ix == nx;
}

}  // namespace

// one (policy, value) slot pair of a model with savings grid a[na]; 0 or -1
int decode_compact_pair(const double* a, int na, const double* rec_c, int n_c, int aux_c, const double* rec_v, int n_v, int aux_v,
                        double* xy_c, double* xy_v) {
    if (!decode_slot(rec_c, n_c, aux_c & 0xffff, (aux_c >> 16) & 0xffff, a, na, nullptr, 0, xy_c)) return -1;
    if (n_v == 0) return (aux_v == 0) ? 0 : -1;
    return decode_slot(rec_v, n_v, aux_v & 0xffff, (aux_v >> 16) & 0xffff, a, na, xy_c, n_c, xy_v) ? 0 : -1;
}

// Expands the compact records of `count` models (slots in layout order, first model = `first` of the batch the tables
// were built for) into densely packed points: offsets[nslots + 1] (points), xy interleaved.  woff: word offsets of the
// slots relative to `words`.  Returns 0, -1 (inconsistent record) or -5 (cap_points too small).
int expand_compact(const HostTables& tab, int nT, int nstate, int first, int count, const double* words, const int64_t* woff,
                   const int32_t* lens, const int32_t* aux, double* xy, size_t cap_points, int64_t* offsets, int nthreads) {
    const size_t per = (size_t)nT * nstate * 4;
    const size_t nsl = per * count;
    offsets[0] = 0;
    for (size_t s = 0; s < nsl; s++) offsets[s + 1] = offsets[s] + lens[s];
    if (!xy) return 0;
    if ((size_t)offsets[nsl] > cap_points) return -5;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > count) nthreads = count > 0 ? count : 1;
    std::vector<int> bad(nthreads, 0);
    auto work = [&](int t) {
        for (int m = t; m < count; m += nthreads) {
            const DevModel& p = tab.models[first + m];
            const double* a = tab.grids.data() + (size_t)p.grid * p.na;
            for (size_t q = 0; q < per; q += 2) {
                const size_t sc = (size_t)m * per + q, sv = sc + 1;
                double* oc = xy + 2 * offsets[sc];
                double* ov = xy + 2 * offsets[sv];
                if (!decode_slot(words + woff[sc], lens[sc], aux[sc] & 0xffff, (aux[sc] >> 16) & 0xffff, a, p.na, nullptr, 0, oc) ||
                    !decode_slot(words + woff[sv], lens[sv], aux[sv] & 0xffff, (aux[sv] >> 16) & 0xffff, a, p.na, oc, lens[sc], ov))
                    bad[t] = 1;
            }
        }
    };
    if (nthreads == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    for (int b : bad)
        if (b) return -1;
    return 0;
}

}  // namespace dcegm
