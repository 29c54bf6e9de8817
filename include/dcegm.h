/*
 * dcegm.h — C ABI of libdcegm_cuda.so, the B200-native DC-EGM solver.
 *
 * This is the drop-in boundary named by BASELINE.json `north_star`: the host
 * side (Julia `ccall`, or the Python/ctypes mirror in dcegm.jl_b200/) fills an
 * array of POD `dcegm_param` (one per model parameterisation) and receives the
 * value/policy polylines the reference stores in `m.v[id,it]` / `m.c[id,it]`
 * (`Envelope.env :: MLine`, a `Vector{Point{Float64}}` — bit-compatible with the
 * interleaved (x,y) doubles used here).
 *
 * The reference has no FFI of its own (pure Julia).  Each entry point below cites
 * the Julia function (file:line under the reference tree) whose work it replaces.
 *
 * Conventions
 *   - every function returns 0 on success or a negative DCEGM_E_* code; nothing
 *     throws or aborts across the ABI; dcegm_last_error() gives the text.
 *   - the caller owns every pointer it passes in and every host destination
 *     buffer; the library owns all device memory and streams (freed by
 *     dcegm_destroy).  One handle per GPU; a handle is used from one host thread
 *     at a time.
 *   - indices are 0-based: it = 0..nT-1 (reference period it+1), id = 0 worker /
 *     1 retiree (reference id = 1 / 2), iy = 0..nstate-1.
 *   - there is NO CPU fallback: without a CUDA device dcegm_create fails.
 */
#ifndef DCEGM_H
#define DCEGM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCEGM_ABI_VERSION 2

/* ---- error codes (function return values) --------------------------------- */
#define DCEGM_OK            0
#define DCEGM_E_INVALID    (-1)  /* bad argument (null pointer, size mismatch, unsupported size) */
#define DCEGM_E_CUDA       (-2)  /* CUDA runtime error, see dcegm_last_error */
#define DCEGM_E_NOSOLUTION (-3)  /* fetch before any solve */
#define DCEGM_E_NODEVICE   (-4)  /* no CUDA device: the library has no CPU path */
#define DCEGM_E_CAPACITY   (-5)  /* destination buffer too small */

/* ---- per-model status bits (dcegm_status) ---------------------------------
 * Each bit corresponds to a condition under which the reference would throw a
 * Julia exception; a failing model never aborts the batch. */
#define DCEGM_ST_OK              0u
#define DCEGM_ST_ISECT_OVERFLOW  (1u << 0) /* more intersections than isec_cap */
#define DCEGM_ST_LINE_OVERFLOW   (1u << 1) /* a line outgrew `cap` points */
#define DCEGM_ST_NONMONOTONE     (1u << 2) /* a stored line of a period > 1 has DECREASING (or unordered) abscissae: the
                                              previous period's interpolate() throws in every Interpolations.jl
                                              version ("knot-vectors must be sorted", line.jl:236) */
#define DCEGM_ST_NO_BRACKET      (1u << 3) /* fzero without sign change (envelope.jl:199,225-228) */
#define DCEGM_ST_MISALIGNED      (1u << 4) /* insert_right shorter than insert_left: insert_right[ileft] is a BoundsError
                                              (dc_algo.jl:486-491) */
#define DCEGM_ST_ALL_FALLING     (1u << 5) /* splitLine returned no section (envelope.jl:97-104) */
#define DCEGM_ST_NONFINITE       (1u << 6) /* NaN/Inf in a computed line */
#define DCEGM_ST_SEG_OVERFLOW    (1u << 7) /* more sections than seg_cap */
#define DCEGM_ST_INSERT_RANGE    (1u << 8) /* kink insertion index out of range (dc_algo.jl:472,489) */
#define DCEGM_ST_LOOP_LIMIT      (1u << 9) /* third-line loop did not terminate (envelope.jl:197-240) */
#define DCEGM_ST_NEG_CONS        (1u << 10) /* simulator: negative consumption, the reference calls error() (simulate.jl:111-116) */
#define DCEGM_ST_YSTATE_RANGE    (1u << 11) /* simulator: Markov draw beyond the last cumulative weight (simulate.jl:148 would index out of bounds) */
#define DCEGM_ST_DUP_KNOT        (1u << 14) /* a stored line of a period > 1 repeats an abscissa (non-decreasing, not strictly
                                               increasing).  The reference's own upper_env produces this whenever an
                                               intersection is found AT a turning point shared by two sections
                                               (envelope.jl:199 f(pt1) == 0, :212 and :248-252 push the point twice).
                                               interpolate() then throws in Interpolations.jl versions that require
                                               unique knots; the lines stored here are what a version without that
                                               check computes (searchsortedlast never selects the empty cell). */
/* Informational bits: the reference returns normally from runf/rung on these models; they neither raise in
 * the host mirror nor remove a model from throughput counts (DCEGM_ST_FATAL_MASK excludes them). */
#define DCEGM_ST_EXTRA_RIGHT     (1u << 12) /* insert_right LONGER than insert_left: the loop of dc_algo.jl:486-492 runs over
                                               insert_left only, the surplus right points are silently dropped */
#define DCEGM_ST_NONMONO_FIRST   (1u << 13) /* a line of the FIRST period is not strictly increasing in x.  The solve never
                                               interpolates period 1 (dc_algo.jl:22-37 reads it+1 only), so runf returns
                                               normally; sim() / plots that interpolate period 1 would throw (line.jl:236) */
#define DCEGM_ST_INFO_MASK       (DCEGM_ST_EXTRA_RIGHT | DCEGM_ST_NONMONO_FIRST)
#define DCEGM_ST_FATAL_MASK      (~DCEGM_ST_INFO_MASK)

/* ---- model parameters: POD mirror of `mutable struct Param` ----------------
 * reference: src/param.jl:33-70, defaults src/param.json, SURVEY.md App. C.
 * Derived fields (neg_gamma, oneminusgamma, ...; param.jl:90-93) are recomputed
 * by the library. */
typedef struct dcegm_param {
    double gamma;        /* CRRA coefficient                    (1.95) */
    double beta;         /* discount factor                     (0.95) */
    double sigma;        /* income shock std                    (0.35) */
    double lambda;       /* taste-shock scale                   (2e-6) */
    double R;            /* gross interest                      (1.0526315789473684) */
    double a_high;       /* upper bound of savings grid         (50) */
    double a_low;        /* lower bound of savings grid         (0) */
    double a_lowT;       /* lower bound in the last period      (0) */
    double cfloor;       /* consumption floor                   (1e-3) */
    double cfloor_plot;  /* unused on this path                 (0.01) */
    double alpha;        /* disutility of work                  (0.35) */
    double alphaT;       /* unused on this path                 (0) */
    double inc0, inc1, inc2; /* income profile                  (0.75, 0.04, 2e-4) */
    double rho;          /* AR1 persistence (GModel)            (0.95) */
    double pension;      /*                                      (0) */
    double delta;        /* prob. of returning to work (FModel) (0) */
    double maxinterest;  /* unused on this path                 (200) */
    double bbar;         /* bequest luxury shifter              (7.7) */
    double nu;           /* bequest strength                    (0) */
    double initw0, initw1; /* simulator initial wealth interval (5, 20) */
    int32_t na;          /* savings grid points                 (100) */
    int32_t ny;          /* income nodes                        (5) */
    int32_t nT;          /* periods                             (25) */
    int32_t nD;          /* discrete choices, must be 2         (2) */
    int32_t k;           /* unused on this path                 (100) */
    int32_t retage;      /* mandatory retirement period         (26) */
    int32_t nsims;       /* simulated agents                    (100) */
    int32_t quad_group;  /* lanes sharing one quadrature sum (power of two, 1 = ascending
                            serial order; no reference field — the reference's order is
                            BLAS-internal, dc_algo.jl:58,74).  0 means 1. */
    int64_t rseed;       /* simulator seed                      (25032020) */
} dcegm_param;

/* fill *p with the defaults of src/param.json */
void dcegm_param_defaults(dcegm_param* p);

/* ---- which solver produced the resident solution -------------------------- */
#define DCEGM_MODEL_F       1   /* dc_EGM!(::FModel)  dc_algo.jl:8-105   */
#define DCEGM_MODEL_G       2   /* dc_EGM!(::GModel)  dc_algo.jl:112-206 */
#define DCEGM_MODEL_MINIMAL 3   /* minimal_EGM        minimal_EGM.jl:3-43 */
#define DCEGM_MODEL_MINIMAL_BORROW  4   /* minimal_EGM_borrow   minimal_EGM.jl:49-89: grid and constraint start at a_low */
#define DCEGM_MODEL_MINIMAL_BEQUEST 5   /* minimal_EGM_bequest  minimal_EGM.jl:92-159: scaleGrid savings grid, bequest
                                           motive nu in the penultimate period, value function kept as (id 0, kind v) */

/* ---- line kinds ------------------------------------------------------------ */
#define DCEGM_KIND_C       0   /* m.c[id,it].env  (policy)            */
#define DCEGM_KIND_V       1   /* m.v[id,it].env  (value)             */
#define DCEGM_KIND_CDIRTY  2   /* m.cdirty[id,it] (dc_algo.jl:79-82)  */
#define DCEGM_KIND_VDIRTY  3   /* m.vdirty[id,it]                     */

/* ---- solve flags ------------------------------------------------------------ */
#define DCEGM_KEEP_DIRTY   (1u << 0)  /* runfdemo: also keep the pre-envelope lines */
#define DCEGM_SUMMARY_ONLY (1u << 1)  /* do not materialise lines in HBM, only summaries */
#define DCEGM_FORCE_FINE_GRID (1u << 2)  /* retirement model: run the fine-grid solver (working lines in global
                                            memory, two launches per period) even when the lines of one model fit into
                                            shared memory.  It is selected automatically for grids that do not
                                            (na > ~1000); results are bit-identical either way. */

/* ---- geometry of the resident solution ------------------------------------- *
 * lines  : double [nmodels][nT][nstate][2 id][2 kind(c,v)][cap][2 (x,y)]
 * lens   : int32  [nmodels][nT][nstate][2 id][2 kind]
 * vbound : double [nmodels][nT][nstate][2 id]           Envelope.vbound (envelope.jl:20)
 * meta   : dcegm_env_meta [nmodels][nT][nstate]         worker (id=0) envelope call
 * isects : double [nmodels][nT][nstate][isec_cap][2]    Envelope.isects
 * removed: uint32 [nmodels][nT][nstate][rem_words]      bit k set <=> index k (0-based, in the
 *          line handed to secondary_envelope, i.e. including bend-back points) is in
 *          Envelope.removed (envelope.jl:64-71)
 * dirty  : double [nmodels][nT][nstate][2 id][2 kind][na][2]   (only with DCEGM_KEEP_DIRTY)
 * For DCEGM_MODEL_MINIMAL nstate = 1 and only (id 0, kind c) is used: x = m[it], y = c[it].
 */
typedef struct dcegm_layout {
    int32_t model;      /* DCEGM_MODEL_* */
    int32_t nmodels;
    int32_t na, ny, nT;
    int32_t nstate;     /* 1 (F, minimal) or ny (G) */
    int32_t cap;        /* points per line slot */
    int32_t isec_cap;
    int32_t rem_words;
    int32_t seg_cap;
    int32_t has_dirty;
    int32_t has_lines;
} dcegm_layout;

typedef struct dcegm_env_meta {
    int32_t n_segments;   /* sections returned by splitLine (envelope.jl:80-126); 1 = no envelope needed */
    int32_t n_isects;     /* length(Envelope.isects) */
    int32_t n_removed;    /* length(Envelope.removed) */
    int32_t n_prepended;  /* bend-back points prepended by do_secondary (dc_algo.jl:406-417) */
} dcegm_env_meta;

/* fixed-size per-model record: the payload of the one allgather of a multi-GPU sweep */
#define DCEGM_SUMMARY_NPROBE 8
typedef struct dcegm_summary {
    int32_t status;            /* DCEGM_ST_* bits */
    int32_t n_env_periods;     /* (it,iy) worker steps where the envelope removed points */
    int32_t tot_segments;
    int32_t tot_isects;
    int32_t tot_removed;
    int32_t max_len;           /* longest stored line */
    int64_t tot_points;        /* sum of all line lengths */
    double  v1_probe[DCEGM_SUMMARY_NPROBE]; /* worker value, period 1, state 0, at probe cash-on-hand */
    double  c1_probe[DCEGM_SUMMARY_NPROBE]; /* worker policy, same points */
    double  checksum_c;        /* sum of all policy ordinates of period 1 */
    double  checksum_v;        /* sum of all finite value ordinates of period 1 */
} dcegm_summary;

typedef struct dcegm_ctx* dcegm_handle;

/* ---- lifetime ---------------------------------------------------------------- */
int  dcegm_abi_version(void);
int  dcegm_create(int device_ordinal, dcegm_handle* out);
int  dcegm_destroy(dcegm_handle h);
const char* dcegm_last_error(dcegm_handle h);   /* valid until the next call on h */

/* pinned host memory for the destination buffers of the fetch calls (optional) */
int  dcegm_host_alloc(void** ptr, size_t bytes);
int  dcegm_host_free(void* ptr);

/* ---- solvers (blocking) -------------------------------------------------------
 * params: host array of nmodels records; all must share na, ny, nT (one launch
 * geometry per call).  The solution stays resident on the device until the next
 * solve or dcegm_destroy.
 *   dcegm_solve_f       replaces runf / runfdemo / dc_EGM!(::FModel)  dc_algo.jl:8-105,335-348
 *                       incl. FModel(p) setup                          param.jl:189-223
 *   dcegm_solve_g       replaces rung / dc_EGM!(::GModel)             dc_algo.jl:112-206,356-362
 *                       incl. GModel(p) setup                          param.jl:261-310
 *   dcegm_solve_minimal replaces minimal_EGM (data, not the plot)     minimal_EGM.jl:3-43
 */
int  dcegm_solve_f(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_solve_g(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_solve_minimal(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
/* the two other teaching models of minimal_EGM.jl; results in the same layout (x = m[it], y = c[it] in
 * (id 0, kind c); minimal_EGM_bequest also y = v[it] in (id 0, kind v)) */
int  dcegm_solve_minimal_borrow(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_solve_minimal_bequest(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);

/* Solve and deliver every stored polyline densely packed into caller-owned host memory in one call
 * (what runf/rung return to the caller: all of m.c / m.v).  xy receives the concatenated (x,y) points
 * of all lines in layout order, offsets[(nmodels*nT*nstate*4)+1] their prefix offsets in points,
 * *npoints the total.  The batch is solved in chunks so that packing and the device-to-host copy of
 * one chunk overlap the kernels of the next; xy from dcegm_host_alloc (pinned) makes the copies
 * asynchronous.  DCEGM_E_CAPACITY if cap_points is too small (nmodels*nT*nstate*4*cap always fits).
 * The solution also stays resident for the fetch calls below. */
int  dcegm_solve_f_packed(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags,
                          double* xy, size_t cap_points, int64_t* offsets, int64_t* npoints);
int  dcegm_solve_g_packed(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags,
                          double* xy, size_t cap_points, int64_t* offsets, int64_t* npoints);

/* ---- results -------------------------------------------------------------------- */
int  dcegm_get_layout(dcegm_handle h, dcegm_layout* out);
/* one polyline; replaces getx/gety(m.c[id,it].env) etc. (line.jl:121-124).  xy receives len
 * interleaved (x,y) pairs; *len is set even when cap_points is too small (then DCEGM_E_CAPACITY). */
int  dcegm_fetch_line(dcegm_handle h, int model, int kind, int id, int iy, int it,
                      double* xy, int cap_points, int* len);
/* bulk copies of models [first, first+count) in the layout documented above */
int  dcegm_fetch_lines(dcegm_handle h, int first, int count, double* lines);
int  dcegm_fetch_lens(dcegm_handle h, int first, int count, int32_t* lens);
int  dcegm_fetch_vbound(dcegm_handle h, int first, int count, double* vbound);
int  dcegm_fetch_meta(dcegm_handle h, int first, int count, dcegm_env_meta* meta);
int  dcegm_fetch_isects(dcegm_handle h, int first, int count, double* isects);
int  dcegm_fetch_removed(dcegm_handle h, int first, int count, uint32_t* removed);
int  dcegm_fetch_dirty(dcegm_handle h, int first, int count, double* dirty);
/* densely packed lines: offsets[(count*nT*nstate*4)+1] prefix offsets in points, xy the
 * concatenated points of all lines in layout order; *npoints total.  Pass xy == NULL to
 * query npoints/offsets only. */
int  dcegm_fetch_packed(dcegm_handle h, int first, int count, double* xy, size_t cap_points,
                        int64_t* offsets, int64_t* npoints);
/* ---- packing without allocation ---------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is dcegm_fetch_packed's
 * device work — prefix offsets of min(lens, cap), then every line slot's points copied behind those of the slots before
 * it — on caller-provided DEVICE memory, with the offsets scratch in a file-scope static device array of the library
 * (csrc/no_alloc.cu) instead of a buffer grown with cudaMalloc.  The call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine); the array (8 MiB + 512 B) is held by every GPU that loads the library until the process
 * exits, and bounds the call: nslots > DCEGM_NOALLOC_MAX_SLOTS is rejected with DCEGM_E_INVALID (a C3 batch of 4096
 * models has 409 600 slots).
 *   lines_dev   double [nslots][cap][2] (the `lines` layout above), lens_dev int32 [nslots]          device, inputs
 *   xy_dev      double [cap_points][2]: the packed points                                            device, output
 *   offsets_dev int64 [nslots + 1] prefix offsets in points, or NULL                                 device, output
 *   npoints     HOST, or NULL.  Non-NULL: the call waits for `stream`, stores the total and returns DCEGM_E_CAPACITY if
 *               it exceeds cap_points.  NULL: the call only enqueues work (one memset, two kernels) and returns.
 *               Either way a slot that would end beyond cap_points is left out, never written past the end.
 *   stream      a cudaStream_t (NULL: the default stream) of device_ordinal, on which all work is enqueued
 *   errbuf      receives the message of a non-zero return (may be NULL)
 * All calls on one device share the array, which keeps the last call's contents (the call clears the words that must
 * start at zero): calls whose GPU work overlaps — different streams, or unordered host threads — race; serialise them. */
#define DCEGM_NOALLOC_MAX_SLOTS (1 << 20)
int  dcegm_pack_lines_noalloc(int device_ordinal, const double* lines_dev, const int32_t* lens_dev, int64_t nslots, int cap,
                              double* xy_dev, size_t cap_points, int64_t* offsets_dev, int64_t* npoints, void* stream,
                              char* errbuf, size_t errbuf_bytes);
/* ---- compact lines: the same polylines in half the bytes --------------------------------------------
 * A stored line is a Vector{Point} (line.jl:1-4), 16 B per point, but nearly every abscissa is redundant: a
 * policy line is (avec + c0, c0) (dc_algo.jl:65), so x = a[i] + y for the savings grid point i a point came from,
 * and the value line of the same (period, state, choice) shares the policy line's abscissae (dc_algo.jl:78, :482,
 * :486-492).  The compact form ships the ordinates plus a few RUNS naming the source of each stretch of abscissae;
 * points for which neither rule reproduces the stored x bit for bit (verified per point on the device) carry an
 * explicit x.  The ordinates, smooth along a line, travel as second or third differences of their 64-bit patterns
 * (integer arithmetic modulo 2^64, exactly invertible), packed in blocks of 8 with a width in four-bit units per block.
 * Record of one line slot (slots in layout order [model][it][iy][id][kind]; an empty slot has no record), in 8-byte words:
 *     head | ysec[yw] | run[nrun] | xs[nx]     head = yw; n = lens[slot], nrun = aux[slot] & 0xffff, nx = aux[slot] >> 16
 *     ysec = I_0 | I_1 - I_0 | control bytes c_b of the ceil(n/8) blocks, zero-padded to 8 | payload, zero-padded to 8
 *       I_k = bit pattern of y_k; R2_k = I_k - 2 I_(k-1) + I_(k-2) for k >= 2 and R2_0 = R2_1 = 0 (place-holders: the
 *       first two ordinates are the raw words in front); R3_k = R2_k - R2_(k-1).  Block b covers k = 8 b .. 8 b + 7
 *       (cnt = 8, or what is left of the line) and holds R2 (c_b bit 5 clear) or R3 (bit 5 set), whichever is narrower,
 *       as two's-complement integers of w = c_b & 31 four-bit units (0: all zero ... 16: raw; bits 6-7 zero): first
 *       cnt x (w >> 1) bytes — the low w >> 1 bytes of residual j, little-endian, at byte j (w >> 1) — then, for odd w,
 *       ceil(cnt / 2) bytes holding the top four bits of residuals 2 i (low half) and 2 i + 1 (high half) in byte i.
 *       Decoding: R2_k = residual, or R2_(k-1) + residual; D_k = D_(k-1) + R2_k with D_1 = I_1 - I_0; I_k = I_(k-1) + D_k.
 *       n = 1: I_0 only
 *     run (64-bit integer) = pos | src << 16 | kind << 32; it covers points pos .. (next run's pos, or n) - 1:
 *       kind 0  x[pos + j] = a[src + j] + y[pos + j]     (a: the model's savings grid, dcegm_get_grid; one IEEE add)
 *       kind 1  x[pos + j] = x of point src + j of the POLICY line of the same (model, it, iy, id)   (value lines only)
 *       kind 2  x[pos + j] = next unread word of xs
 * woffsets[s] is the word offset of slot s's record in `words`, woffsets[nslots] the total (*nwords).  IDENTICAL LINES
 * TRAVEL ONCE: a (policy, value) slot pair whose two lines are bit-identical to the same (it, iy, id) pair of the FIRST
 * model of its delivery chunk (dcegm_fetch_compact: of model `first`) has no record of its own — its woffsets entries
 * are that model's, its aux / lens equal that model's — so woffsets is not monotone, and the size of a record follows
 * from lens, aux and the record's head, never from differences of woffsets.  In a sweep over (lambda, sigma) all retiree
 * lines are of this kind (a retiree has no income draw, param.jl:212, and no taste-shocked choice, dc_algo.jl:44): 44 % of a C3
 * batch's bytes.  Decoders need no special case.  DCEGM_COMPACT_ALIAS=0 (environment) switches the detection off.
 * dcegm_solve_*_compact is dcegm_solve_*_packed with
 * this delivery format (same chunked pipeline; `words` from dcegm_host_alloc makes the copies asynchronous);
 * cap_words = nslots * (2 * cap + cap / 64 + 8) always fits (a C3 batch needs 0.24 words per point).  dcegm_fetch_compact encodes models [first, first + count) of the
 * resident solution (words == NULL: sizes only).  dcegm_expand_compact rebuilds the densely packed points of
 * dcegm_fetch_packed (bit-identical: tests/test_gpu_compact.py) on `nthreads` host threads, for the batch of the
 * handle's last solve; woffsets / lens / aux point at the first slot of model `first`, woffsets values are relative
 * to `words`.  dcegm_decode_compact_pair is the handle-free primitive a lazy host mirror calls per Envelope: the
 * policy record (and optionally the value record) of one (model, it, iy, id) -> interleaved points. */
int  dcegm_solve_f_compact(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags, double* words,
                           size_t cap_words, int64_t* woffsets, int32_t* lens, int32_t* aux, int64_t* nwords);
int  dcegm_solve_g_compact(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags, double* words,
                           size_t cap_words, int64_t* woffsets, int32_t* lens, int32_t* aux, int64_t* nwords);
int  dcegm_fetch_compact(dcegm_handle h, int first, int count, double* words, size_t cap_words, int64_t* woffsets,
                         int32_t* lens, int32_t* aux, int64_t* nwords);
int  dcegm_expand_compact(dcegm_handle h, int first, int count, const double* words, const int64_t* woffsets,
                          const int32_t* lens, const int32_t* aux, double* xy, size_t cap_points, int64_t* offsets,
                          int64_t* npoints, int nthreads);
int  dcegm_decode_compact_pair(const double* grid, int na, const double* rec_c, int n_c, int aux_c, const double* rec_v,
                               int n_v, int aux_v, double* xy_c, double* xy_v);
/* the savings grid avec of one model of the last uploaded batch (param.jl:199, :288); grid == NULL: *na only */
int  dcegm_get_grid(dcegm_handle h, int model, double* grid, int cap, int* na);
int  dcegm_status(dcegm_handle h, int32_t* per_model_status);   /* nmodels entries */
int  dcegm_summaries(dcegm_handle h, dcegm_summary* dst);       /* nmodels records, host */
/* device-to-device copy of the summaries into caller-provided device memory (the send
 * buffer of the NCCL allgather of a multi-GPU sweep) */
int  dcegm_summaries_to_device(dcegm_handle h, void* dst_device, size_t bytes);

/* ---- timing of the last solve, measured with CUDA events on the library's stream ---- */
typedef struct dcegm_timing {
    float  setup_host_ms;   /* FModel/GModel grids and tables on the host */
    float  h2d_ms;
    float  kernel_ms;       /* all solver kernels */
    int32_t n_launches;     /* kernels launched by the solve */
    int32_t _pad;
} dcegm_timing;
int  dcegm_last_timing(dcegm_handle h, dcegm_timing* out);

/* ---- split solve for benchmarking with inputs already resident in HBM --------------
 * dcegm_upload_f does the host setup + H2D; dcegm_run launches the solver kernels on the
 * data already resident (can be called repeatedly). */
int  dcegm_upload_f(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_upload_g(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_upload_minimal(dcegm_handle h, const dcegm_param* params, int nmodels, unsigned flags);
int  dcegm_run(dcegm_handle h);

/* ---- forward simulation of the resident solution ------------------------------------------
 * Replaces sim(m::FModel,p) (simulate.jl:61-130) and sim(m::GModel,p) (simulate.jl:133-203) for every
 * model of the resident batch: p.nsims agents per model, initial wealth U(initw0,initw1), policies
 * and choice probabilities interpolated from the stored polylines (interp, vfun, ccp as in the solver).
 *
 * Shocks.  shocks == NULL: counter-based Philox4x32-10 draws keyed by p.rseed with counter (agent,
 * period, stream) — see csrc/dc_rng.h; like the reference's Random.seed!(p.rseed) the model index does
 * not enter.  Otherwise host arrays of the reference's raw rand() draws, column-major [agent + nsims*it]
 * (all models then share nsims): wshocks (FModel: pushed through the normal quantile and scaled by
 * sigma, simulate.jl:73; GModel: Markov transition draws, simulate.jl:138,148), dshocks, rshocks
 * (FModel only, may be NULL when delta == 0), w0shocks [nsims], ystate0 [nsims] (GModel, 1-based).
 *
 * Results.  Always: per (model, period) means (moments) and the histogram of retirement ages.
 * With DCEGM_SIM_PANEL (all models must share nsims) the full Simulation panel (simulate.jl:1-28) is
 * kept on the device for dcegm_fetch_sim_panel.
 */
#define DCEGM_SIM_PANEL (1u << 0)
#define DCEGM_SIM_NMOM 8
/* moment k of period it: mean over agents of  0 w0, 1 w1, 2 cons, 3 util, 4 inc, 5 prob_work,
 * 6 share with worker == 1, 7 share with negative consumption.  Sums run over tiles of 1024 agents
 * (4 per thread of a 256-thread block, ascending), each tile reduced by the fixed shuffle tree, tiles
 * added in ascending order — the CPU oracle reproduces the same order bit for bit. */
typedef struct dcegm_sim_shocks {
    const double* wshocks;
    const double* dshocks;
    const double* rshocks;
    const double* w0shocks;
    const int32_t* ystate0;
    int32_t nsims; /* rows of the matrices above */
    int32_t _pad;
} dcegm_sim_shocks;
int  dcegm_simulate(dcegm_handle h, unsigned flags, const dcegm_sim_shocks* shocks);
/* moments: double [count][nT][DCEGM_SIM_NMOM]; ret_hist: int32 [count][nT+1], entry a = agents with
 * ret_age == a (0 = never retired; simulate.jl:95); sim_status: int32 [count] DCEGM_ST_* bits.  Any may be NULL. */
int  dcegm_fetch_sim_moments(dcegm_handle h, int first, int count, double* moments, int32_t* ret_hist,
                             int32_t* sim_status);
/* device-to-device copy of the moments (nmodels*nT*DCEGM_SIM_NMOM doubles) into caller-provided device
 * memory: the second part of the allgather payload of a multi-GPU sweep */
int  dcegm_sim_moments_to_device(dcegm_handle h, void* dst_device, size_t bytes);
/* The payload of the ONE all-gather of a multi-GPU sweep: one fixed-size record per model of the
 * resident batch, written to caller-provided DEVICE memory (the NCCL send buffer):
 *   [ dcegm_summary | int32 sim_status (-1: not simulated) | int32 nT | double moments[nT_pad][DCEGM_SIM_NMOM] ]
 * nT_pad >= nT lets batches of different nT share one record size (a sweep over grid sizes runs one solve per
 * (na, ny, nT) group).  Moments are NaN beyond nT, when dcegm_simulate has not run, and for models whose status has a
 * fatal bit or DCEGM_ST_NONMONO_FIRST (sim() interpolates period 1): a flagged model's moments are not a result. */
size_t dcegm_sweep_record_bytes(int nT_pad);
int  dcegm_sweep_records_to_device(dcegm_handle h, void* dst_device, size_t bytes, int nT_pad);
/* one model's panel, column-major [agent + nsims*it] like the reference's matrices; ret_age [nsims].
 * Any destination may be NULL. */
int  dcegm_fetch_sim_panel(dcegm_handle h, int model, double* w0, double* w1, double* cons, double* util,
                           double* inc, double* prob_work, int32_t* ystate, int32_t* worker, int32_t* ret_age);

/* ---- diagnostics ---------------------------------------------------------------------
 * Runs n random operand pairs with binary exponents in [emin, emax] (plus exact zeros and exact
 * quotients) through the kernels' branch-free division / reciprocal (csrc/dc_math.h) and counts
 * results that differ from the IEEE operations: mismatches2[0] divisions, [1] reciprocals. */
int  dcegm_debug_math_check(unsigned long long n, unsigned long long seed, int emin, int emax,
                            unsigned long long* mismatches2);

/* Every device buffer of a handle ends in a 4 KB guard band written at allocation; *nbad counts the buffers whose band
 * was overwritten since (a kernel wrote past the end of the buffer), *nbuffers the buffers checked. */
int  dcegm_debug_check_guards(dcegm_handle h, int* nbuffers, int* nbad);

/* The host-side model constructors of the library (csrc/host_setup.cpp), callable without a device, so that the
 * tables the kernels receive can be checked against independent implementations (tests/test_host_setup_tables.py):
 * which = 0 range(p0, stop=p1, length=n) -> out0[n] (param.jl:199);  1 quadpoints(n,0,1) -> nodes out0[n], weights
 * out1[n] (utils.jl:261-297);  2 quantile(Normal(0,1), p0) -> out0[0] (param.jl:195-196);  3 gausshermite(n) -> out0[n],
 * out1[n] (param.jl:279);  4 rouwenhorst(rho=p0, mu=p1, sigma=p2, n) -> support out0[n], transition matrix out1[n*n]
 * row-major (utils.jl:178-197);  5 scaleGrid(p0, p1, n; logorder=iarg) -> out0[n] (utils.jl:200-238). */
int  dcegm_debug_host_table(int which, int n, double p0, double p1, double p2, int iarg, double* out0, double* out1);

/* FP64 issue rate of the device: thread-level DFMA per second retired by a kernel of independent DFMA chains
 * at full occupancy (no memory traffic); *ms = duration of the best repetition.  The solver kernels are bound by
 * this pipe, not by HBM ("Roofline"); bench.py reports their FP64 instruction rate against it. */
int  dcegm_debug_dfma_rate(int device_ordinal, double* dfma_per_s, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* DCEGM_H */
