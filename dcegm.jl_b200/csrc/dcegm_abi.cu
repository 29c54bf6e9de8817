// dcegm_abi.cu — the C ABI of libdcegm_cuda.so (include/dcegm.h): handle, device memory, streams,
// uploads, kernel launches and result transfers.  No solver arithmetic lives here and there is no
// CPU path: every solve runs the sm_100a kernels of solver_*.cu.
#include <cuda.h>  // types of the stream memory operation fetched through cudaGetDriverEntryPoint (libcuda is not linked)
#include <cuda_runtime.h>

#include <sys/syscall.h>
#include <unistd.h>

#include <cctype>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/dcegm.h"
#include "device_common.cuh"
#include "host_setup.hpp"

namespace dcegm {
cudaError_t launch_solve_f(const DeviceBatch& B, int first, int count, cudaStream_t stream);
cudaError_t launch_solve_f_big(const DeviceBatch& B, int first, int count, unsigned char* scratch, cudaStream_t stream,
                               int* nlaunch);
size_t solver_f_big_scratch_bytes(int cap, int isec_cap);
cudaError_t launch_solve_g_big(const DeviceBatch& B, int first, int count, unsigned char* scratch, cudaStream_t stream,
                               int* nlaunch);
size_t solver_g_big_scratch_bytes(int cap, int isec_cap, int ny);
cudaError_t launch_solve_g(const DeviceBatch& B, int first, int count, cudaStream_t stream);
cudaError_t launch_solve_minimal(const DeviceBatch& B, int first, int count, cudaStream_t stream);
cudaError_t launch_pack_lines(const DeviceBatch& B, int first, int count, const long long* d_offsets, double2* d_dst,
                              cudaStream_t stream);
cudaError_t launch_chunk_offsets(const int32_t* lens, size_t nsl, int cap, long long* d_off, long long* h_off_mapped,
                                 cudaStream_t stream);
cudaError_t launch_compact_count(const DeviceBatch& B, int first, int count, bool alias, int32_t* aux, int32_t* yws, int32_t* wsize, unsigned* codes,
                                 int32_t* aux_host, int32_t* lens_host, cudaStream_t stream);
cudaError_t launch_compact_write(const DeviceBatch& B, int first, int count, const int32_t* aux, const int32_t* yws, const unsigned* codes,
                                 const long long* off, double* dst, cudaStream_t stream);
cudaError_t launch_sweep_records(const dcegm_summary* summ, const double* mom, const int32_t* sim_status, int nmodels, int nT,
                                 int nT_pad, size_t rec_bytes, unsigned char* dst, cudaStream_t stream);
cudaError_t launch_sim(const DeviceBatch& B, const SimBatch& Sb, int first, int count, int max_smem_optin, cudaStream_t stream);
size_t sim_smem_bytes(const dcegm_layout& l, bool global_lines);
int expand_compact(const HostTables& tab, int nT, int nstate, int first, int count, const double* words, const int64_t* woff,
                   const int32_t* lens, const int32_t* aux, double* xy, size_t cap_points, int64_t* offsets, int nthreads);
int decode_compact_pair(const double* a, int na, const double* rec_c, int n_c, int aux_c, const double* rec_v, int n_v, int aux_v,
                        double* xy_c, double* xy_v);
size_t solver_f_smem_bytes(int cap, int isec_cap);
size_t solver_g_smem_bytes(int cap, int isec_cap, int na, int ny);
}  // namespace dcegm

using namespace dcegm;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct dcegm_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t s_pack = nullptr, s_copy = nullptr;  // pipelined solve-to-host: packing and D2H streams
    cudaStream_t s_pack_lo = nullptr, s_pack_hi = nullptr;  // ... the packing stream at default / highest priority
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::vector<cudaEvent_t> ev_pool;                 // per-chunk events of the pipelined solve
    long long* pin_off = nullptr;
    int32_t* pin_aux = nullptr;   // compact delivery: per-slot run/explicit counts and lengths of a chunk (host-mapped)
    size_t pin_slots = 0;
    DevBuf cmp_aux, cmp_wsize, cmp_codes, cmp_yw;
    DevBuf order;         // launch order of whole-batch solves (costliest models first)
    DevBuf chunk_done, model_chunk;  // one-launch pipelined solve: per-chunk completion counters, chunk of every model
    int wait_value_state = 0;        // 0 untested, 1 cuStreamWaitValue32 works, -1 unavailable (chunked launches instead)
    // host tables of the last upload, page-locked in place (cudaHostRegister) so that their H2D copies run at full
    // PCIe rate instead of through the driver's staging buffer; re-registered only when a vector moved or grew
    struct Pinned { const void* p = nullptr; size_t bytes = 0; };
    Pinned pin_tab[6];
    long long pin_sig[5] = {-1, -1, -1, -1, -1};  // (model, nmodels, na, ny, nT) the registrations belong to
    DevBuf pack_xy2;
    DevBuf big_scratch;   // fine-grid retirement solver (solver_f_big.cu): per-model working lines + envelope scratch
    bool big = false;     // the uploaded batch runs on the fine-grid solver
    int launches_last = 0;
    DevBuf simm, yvec, mom, rhist, sstat, pan_d, pan_i, pan_ret, sh_w, sh_d, sh_r, sh_w0, sh_y0;
    bool simulated = false;
    int sim_nsims = 0;      // panel geometry of the last simulation (0 = no panel kept)
    float sim_ms = 0.f;
    std::string err;
    DevBuf models, grids, off, wgt, lines, lens, vbound, meta, isects, removed, dirty, status, summaries, pack_off, pack_xy;
    DeviceBatch batch{};
    HostTables tab;
    bool uploaded = false, solved = false;
    unsigned flags = 0;
    int max_smem_optin = 0;
    dcegm_timing timing{};
};

namespace {

int fail(dcegm_ctx* h, int code, const std::string& msg) {
    if (h) h->err = msg;
    return code;
}
int cuda_fail(dcegm_ctx* h, cudaError_t e, const char* what) {
    return fail(h, DCEGM_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                                \
    do {                                                        \
        cudaError_t e_ = (call);                                \
        if (e_ != cudaSuccess) return cuda_fail(h, e_, #call);  \
    } while (0)

// Every device buffer ends in a guard band of kGuardBytes filled with 0xA5 when it is allocated: a kernel that writes
// past the end of a buffer changes it, and dcegm_debug_check_guards() finds out (compute-sanitizer is not available on
// this GPU pool; tests/test_gpu_guards.py runs every solver and checks the bands).
constexpr size_t kGuardBytes = 4096;
int ensure(dcegm_ctx* h, DevBuf& b, size_t bytes, bool zero_new) {
    if (bytes <= b.cap) return DCEGM_OK;
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    CK(cudaMalloc(&b.p, bytes + kGuardBytes));
    b.cap = bytes;
    if (zero_new) CK(cudaMemsetAsync(b.p, 0, bytes, h->stream));
    CK(cudaMemsetAsync((char*)b.p + bytes, 0xA5, kGuardBytes, h->stream));
    return DCEGM_OK;
}

int upload(dcegm_ctx* h, int model, const dcegm_param* params, int nmodels, unsigned flags) {
    if (!h) return DCEGM_E_INVALID;
    if (!params || nmodels <= 0) return fail(h, DCEGM_E_INVALID, "params is null or nmodels <= 0");
    CK(cudaSetDevice(h->device));
    h->uploaded = h->solved = false;
    {
        // a batch of another geometry reallocates the table vectors: release their page locks first
        const long long sig[5] = {model, nmodels, params[0].na, params[0].ny, params[0].nT};
        if (memcmp(sig, h->pin_sig, sizeof sig) != 0) {
            for (auto& r : h->pin_tab) {
                if (r.p) cudaHostUnregister(const_cast<void*>(r.p));
                r.p = nullptr;
                r.bytes = 0;
            }
            memcpy(h->pin_sig, sig, sizeof sig);
        }
    }
    auto t0 = std::chrono::steady_clock::now();
    const char* why = "";
    if (!build_tables(model, params, nmodels, h->tab, &why)) return fail(h, DCEGM_E_INVALID, why);
    auto t1 = std::chrono::steady_clock::now();
    h->timing.setup_host_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();

    const int na = params[0].na, ny = params[0].ny, nT = params[0].nT;
    dcegm_layout lay;
    make_layout(model, nmodels, na, ny, nT, (flags & DCEGM_KEEP_DIRTY) ? 1 : 0, (flags & DCEGM_SUMMARY_ONLY) ? 0 : 1, &lay);
    if (model == DCEGM_MODEL_F && ny > kMaxNy) return fail(h, DCEGM_E_INVALID, "ny > 64 is not supported by the retirement-model solver");
    size_t smem = model == DCEGM_MODEL_F   ? solver_f_smem_bytes(lay.cap, lay.isec_cap)
                  : model == DCEGM_MODEL_G ? solver_g_smem_bytes(lay.cap, lay.isec_cap, na, ny)
                                           : (size_t)(6 * (na + 1) * 8 + 1024);
    // retirement / general model on a grid too fine for one SM's shared memory: the fine-grid solvers (lines in
    // global memory, solver_f_big.cu / solver_g_big.cu)
    const bool can_big = model == DCEGM_MODEL_F || model == DCEGM_MODEL_G;
    h->big = can_big &&
             ((flags & DCEGM_FORCE_FINE_GRID) || lay.cap >= 2 * kSearchTop || (long long)smem > (long long)h->max_smem_optin);
    if (h->big) lay.seg_cap = lay.cap;
    if (h->big && model == DCEGM_MODEL_G && ny > kMaxNy) return fail(h, DCEGM_E_INVALID, "ny > 64 is not supported by the fine-grid general-model solver");
    if (h->big && lay.cap >= 65536) return fail(h, DCEGM_E_INVALID, "na too large: polylines are indexed with 16 bits (na < ~52000)");
    if (!h->big && lay.cap >= 2 * kSearchTop) return fail(h, DCEGM_E_INVALID, "na too large for the shared-memory resident solver");
    if (!h->big && (long long)smem > (long long)h->max_smem_optin) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "na=%d needs %zu B of shared memory per CTA (device limit %d B): grid too fine for the "
                 "shared-memory resident solver",
                 na, smem, h->max_smem_optin);
        return fail(h, DCEGM_E_INVALID, msg);
    }

    const size_t nslots = (size_t)nmodels * nT * lay.nstate * 4;
    const size_t nstates = (size_t)nmodels * nT * lay.nstate;
    int rc;
    if ((rc = ensure(h, h->models, sizeof(DevModel) * nmodels, false))) return rc;
    if ((rc = ensure(h, h->grids, sizeof(double) * h->tab.grids.size(), false))) return rc;
    if ((rc = ensure(h, h->off, sizeof(double) * h->tab.off.size(), false))) return rc;
    if ((rc = ensure(h, h->wgt, sizeof(double) * h->tab.wgt.size(), false))) return rc;
    if (lay.has_lines)
        if ((rc = ensure(h, h->lines, sizeof(double2) * nslots * lay.cap, true))) return rc;
    if ((rc = ensure(h, h->lens, sizeof(int32_t) * nslots, true))) return rc;
    if ((rc = ensure(h, h->vbound, sizeof(double) * nstates * 2, true))) return rc;
    if ((rc = ensure(h, h->meta, sizeof(dcegm_env_meta) * nstates, true))) return rc;
    if ((rc = ensure(h, h->isects, sizeof(double2) * nstates * lay.isec_cap, true))) return rc;
    if ((rc = ensure(h, h->removed, sizeof(uint32_t) * nstates * lay.rem_words, true))) return rc;
    if (lay.has_dirty)
        if ((rc = ensure(h, h->dirty, sizeof(double2) * nslots * na, true))) return rc;
    if ((rc = ensure(h, h->status, sizeof(int32_t) * nmodels, true))) return rc;
    if ((rc = ensure(h, h->summaries, sizeof(dcegm_summary) * nmodels, true))) return rc;
    if (h->big) {
        const size_t per_model = model == DCEGM_MODEL_F ? solver_f_big_scratch_bytes(lay.cap, lay.isec_cap)
                                                        : solver_g_big_scratch_bytes(lay.cap, lay.isec_cap, ny);
        if ((rc = ensure(h, h->big_scratch, per_model * nmodels, false))) return rc;
    }

    {
        const void* ptr[6] = {h->tab.models.data(), h->tab.grids.data(), h->tab.off.data(), h->tab.wgt.data(), h->tab.sim.data(), h->tab.yvec.data()};
        const size_t len[6] = {sizeof(DevModel) * h->tab.models.size(), 0 /* grids: size varies with the parameters */, sizeof(double) * h->tab.off.size(),
                               sizeof(double) * h->tab.wgt.size(), sizeof(SimModel) * h->tab.sim.size(), sizeof(double) * h->tab.yvec.size()};
        for (int i = 0; i < 6; i++) {
            dcegm_ctx::Pinned& r = h->pin_tab[i];
            if (r.p == ptr[i] && r.bytes == len[i]) continue;
            if (r.p) cudaHostUnregister(const_cast<void*>(r.p));
            r.p = nullptr;
            r.bytes = 0;
            if (len[i] >= (64u << 10) && cudaHostRegister(const_cast<void*>(ptr[i]), len[i], cudaHostRegisterDefault) == cudaSuccess) {
                r.p = ptr[i];
                r.bytes = len[i];
            } else {
                cudaGetLastError();  // small tables, or registration refused: pageable copy
            }
        }
    }
    CK(cudaEventRecord(h->ev[0], h->stream));
    CK(cudaMemcpyAsync(h->models.p, h->tab.models.data(), sizeof(DevModel) * nmodels, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->grids.p, h->tab.grids.data(), sizeof(double) * h->tab.grids.size(), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->off.p, h->tab.off.data(), sizeof(double) * h->tab.off.size(), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->wgt.p, h->tab.wgt.data(), sizeof(double) * h->tab.wgt.size(), cudaMemcpyHostToDevice, h->stream));
    if ((rc = ensure(h, h->simm, sizeof(SimModel) * nmodels, false))) return rc;
    if ((rc = ensure(h, h->yvec, sizeof(double) * h->tab.yvec.size(), false))) return rc;
    CK(cudaMemcpyAsync(h->simm.p, h->tab.sim.data(), sizeof(SimModel) * nmodels, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->yvec.p, h->tab.yvec.data(), sizeof(double) * h->tab.yvec.size(), cudaMemcpyHostToDevice, h->stream));
    CK(cudaEventRecord(h->ev[1], h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaEventElapsedTime(&h->timing.h2d_ms, h->ev[0], h->ev[1]));
    h->simulated = false;

    DeviceBatch& B = h->batch;
    B.lay = lay;
    B.models = (const DevModel*)h->models.p;
    B.grids = (const double*)h->grids.p;
    B.off = (const double*)h->off.p;
    B.wgt = (const double*)h->wgt.p;
    B.lines = lay.has_lines ? (double2*)h->lines.p : nullptr;
    B.lens = (int32_t*)h->lens.p;
    B.vbound = (double*)h->vbound.p;
    B.meta = (dcegm_env_meta*)h->meta.p;
    B.isects = (double2*)h->isects.p;
    B.removed = (uint32_t*)h->removed.p;
    B.dirty = lay.has_dirty ? (double2*)h->dirty.p : nullptr;
    B.status = (int32_t*)h->status.p;
    B.summaries = (dcegm_summary*)h->summaries.p;
    B.order = nullptr;
    B.chunk_done = nullptr;
    B.model_chunk = nullptr;
    // Whole-batch solves start the costliest models first (longest-processing-time order): one CTA solves one model, and
    // the envelope makes small-lambda models several times dearer than smooth ones, so the last wave should be cheap ones.
    // The taste-shock scale is the predictor.  Chunked solves (dcegm_solve_*_packed/_compact) keep the batch order.
    {
        const char* e = getenv("DCEGM_LPT");
        if ((model == DCEGM_MODEL_F || model == DCEGM_MODEL_G) && !h->big && nmodels > 1 && (!e || atoi(e) != 0)) {
            std::vector<int32_t> ord(nmodels);
            for (int i = 0; i < nmodels; i++) ord[i] = i;
            std::stable_sort(ord.begin(), ord.end(), [&](int32_t a, int32_t b) { return params[a].lambda < params[b].lambda; });
            if ((rc = ensure(h, h->order, sizeof(int32_t) * nmodels, false))) return rc;
            CK(cudaMemcpyAsync(h->order.p, ord.data(), sizeof(int32_t) * nmodels, cudaMemcpyHostToDevice, h->stream));
            CK(cudaStreamSynchronize(h->stream));
            B.order = (const int32_t*)h->order.p;
        }
    }
    h->flags = flags;
    h->uploaded = true;
    return DCEGM_OK;
}

cudaError_t launch_range(dcegm_ctx* h, const DeviceBatch& B, int first, int count, cudaStream_t stream) {
    h->launches_last = 1;
    if (h->big && B.lay.model == DCEGM_MODEL_G)
        return launch_solve_g_big(B, first, count, (unsigned char*)h->big_scratch.p, stream, &h->launches_last);
    if (h->big) return launch_solve_f_big(B, first, count, (unsigned char*)h->big_scratch.p, stream, &h->launches_last);
    if (B.lay.model == DCEGM_MODEL_F) return launch_solve_f(B, first, count, stream);
    if (B.lay.model == DCEGM_MODEL_G) return launch_solve_g(B, first, count, stream);
    return launch_solve_minimal(B, first, count, stream);
}

int run(dcegm_ctx* h) {
    if (!h) return DCEGM_E_INVALID;
    if (!h->uploaded) return fail(h, DCEGM_E_NOSOLUTION, "dcegm_run before dcegm_upload_*");
    CK(cudaSetDevice(h->device));
    const DeviceBatch& B = h->batch;
    const size_t nstates = (size_t)B.lay.nmodels * B.lay.nT * B.lay.nstate;
    CK(cudaEventRecord(h->ev[2], h->stream));
    // terminal-period slots of the per-state side tables are not written by the solver
    CK(cudaMemsetAsync(B.isects, 0, sizeof(double2) * nstates * B.lay.isec_cap, h->stream));
    CK(cudaMemsetAsync(B.removed, 0, sizeof(uint32_t) * nstates * B.lay.rem_words, h->stream));
    CK(cudaMemsetAsync(B.lens, 0, sizeof(int32_t) * nstates * 4, h->stream));
    CK(cudaMemsetAsync(B.meta, 0, sizeof(dcegm_env_meta) * nstates, h->stream));
    CK(cudaMemsetAsync(B.vbound, 0, sizeof(double) * nstates * 2, h->stream));
    cudaError_t e = launch_range(h, B, 0, B.lay.nmodels, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "kernel launch");
    CK(cudaEventRecord(h->ev[3], h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaEventElapsedTime(&h->timing.kernel_ms, h->ev[2], h->ev[3]));
    h->timing.n_launches = h->launches_last;
    h->solved = true;
    return DCEGM_OK;
}

// Compact delivery: slot pairs whose lines are bit-identical to the same pair of the first model of their chunk carry
// no record (compact_count_kernel marks them aux = -1).  off[0 .. nsl] / aux[0 .. nsl): the chunk's word offsets and run
// counts, `per` slots per model; marked slots take over the first model's offset and run counts.  DCEGM_COMPACT_ALIAS=0
// switches the detection off (every slot then has its own record).
bool compact_alias() {
    const char* e = getenv("DCEGM_COMPACT_ALIAS");
    return !(e && atoi(e) == 0);
}
void resolve_aliases(int64_t* off, int32_t* aux, size_t per, size_t nsl) {
    for (size_t s = per; s < nsl; s++)
        if (aux[s] == -1) {
            off[s] = off[s % per];
            aux[s] = aux[s % per];
        }
}

// Pipelined solve-to-host: the batch is solved in chunks on the compute stream while finished chunks
// are packed densely (s_pack) and copied to the caller's buffer (s_copy), so that the D2H transfer of
// chunk c overlaps the kernels of chunk c+1...  The only host round trip per chunk is its line lengths
// (for the prefix offsets the caller gets back anyway).
// compact != 0: the chunks are delivered in the compact line format (compact_lines.cu; include/dcegm.h): xy receives
// 8-byte words, cap_points / offsets / npoints count words, lens_out / aux_out receive the per-slot lengths and
// run counts.
int solve_packed(dcegm_ctx* h, int model, const dcegm_param* params, int nmodels, unsigned flags, double* xy,
                 size_t cap_points, int64_t* offsets, int64_t* npoints, int compact = 0, int32_t* lens_out = nullptr,
                 int32_t* aux_out = nullptr) {
    if (!h) return DCEGM_E_INVALID;
    if (!xy || !offsets || !npoints) return fail(h, DCEGM_E_INVALID, "xy/offsets/npoints is null");
    if (compact && (!lens_out || !aux_out)) return fail(h, DCEGM_E_INVALID, "lens/aux is null");
    if (flags & DCEGM_SUMMARY_ONLY) return fail(h, DCEGM_E_INVALID, "DCEGM_SUMMARY_ONLY keeps no lines to pack");
    const bool trace = getenv("DCEGM_TRACE") != nullptr;  // host timestamps of the pipeline on stderr
    const auto tr0 = std::chrono::steady_clock::now();
    auto trace_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tr0).count(); };
    int rc = upload(h, model, params, nmodels, flags);
    if (rc) return rc;
    if (trace) fprintf(stderr, "[dcegm] %8.3f ms upload done (host setup %.3f, h2d %.3f)\n", trace_ms(), h->timing.setup_host_ms, h->timing.h2d_ms);
    const DeviceBatch& B = h->batch;
    const dcegm_layout& l = B.lay;
    const size_t per = (size_t)l.nT * l.nstate * 4;
    const size_t nstates = (size_t)nmodels * l.nT * l.nstate;
    // chunks of whole waves (2 CTAs x 148 SMs): one wave first, so that the first transfer starts early, then
    // two waves each, at most 16 chunks
    // ONE launch for the whole batch when the driver offers stream memory operations: finished models count up their
    // chunk's counter and the packing stream (highest priority: its CTAs take the SM slots solver CTAs free) waits for
    // the counter instead of for a per-chunk kernel — the GPU never drains at a chunk boundary, and chunks can be small.
    // DCEGM_ONE_LAUNCH=0 keeps one launch per chunk (also the fine-grid solvers' way).
    typedef CUresult (*wait_value_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
    static wait_value_fn wait_value = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
        cudaGetLastError();
        return (wait_value_fn)fn;
    }();
    const char* ol = getenv("DCEGM_ONE_LAUNCH");
    const bool want_one = !h->big && (model == DCEGM_MODEL_F || model == DCEGM_MODEL_G) && wait_value && h->wait_value_state >= 0 &&
                          !(ol && atoi(ol) == 0) && nmodels > 296;
    const int wave = 296;
    std::vector<int> cstart, ccount;
    {
        int per_chunk = want_one ? wave : 2 * wave;
        while ((nmodels - wave + per_chunk - 1) / per_chunk + 2 > (want_one ? 32 : 16)) per_chunk += wave;
        int c0 = 0;
        while (c0 < nmodels) {
            // half a wave first (one CTA per SM: the shortest time to the first transfer), then one wave, then per_chunk
            int cnt = nmodels > 2 * wave ? (cstart.empty() ? wave / 2 : cstart.size() == 1 ? wave : per_chunk) : per_chunk;
            if (want_one && nmodels - c0 <= per_chunk + wave / 2 && nmodels - c0 > wave / 2) cnt = nmodels - c0 - wave / 2;  // short last chunk
            // fine-grid solvers: one envelope CTA per SM and chunk.  (Splitting the last chunk to shorten the one transfer
            // nothing can hide was measured slower — 148 + 65 + 43 models: 426 ms against 401 ms for 148 + 108: small
            // chunks solve inefficiently, and the compact records of 108 models are a 26 ms copy.)
            if (h->big) cnt = 148 * ((nmodels + 148 * 16 - 1) / (148 * 16));
            cnt = cnt < nmodels - c0 ? cnt : nmodels - c0;
            cstart.push_back(c0);
            ccount.push_back(cnt);
            c0 += cnt;
        }
    }
    const int nchunks = (int)cstart.size();
    int chunk = 0;
    for (int c : ccount) chunk = c > chunk ? c : chunk;
    if (!h->s_pack_lo) {
        int prio_lo = 0, prio_hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        CK(cudaStreamCreateWithPriority(&h->s_pack_lo, cudaStreamNonBlocking, prio_lo));
        CK(cudaStreamCreateWithPriority(&h->s_pack_hi, cudaStreamNonBlocking, prio_hi));
        CK(cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking));
    }
    {
        // Priority of the packing kernels.  One launch: highest, or they would wait for the solver grid's last pending
        // CTA (measured: first chunk encoded at 18.8 ms instead of 5.3 ms).  One launch per chunk: default — at the
        // highest priority they take SM slots from the solver, which then feeds the copy engine late (133.7 k vs
        // 116.4 k models/s end to end).  DCEGM_PACK_PRIO=0/1 overrides.
        const char* pp = getenv("DCEGM_PACK_PRIO");
        const bool hi = pp ? atoi(pp) != 0 : want_one;
        h->s_pack = hi ? h->s_pack_hi : h->s_pack_lo;
    }
    while (h->ev_pool.size() < (size_t)3 * nchunks) {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_pool.push_back(e);
    }
    const size_t cslots = per * chunk;
    if (h->pin_slots < cslots) {
        if (h->pin_off) cudaFreeHost(h->pin_off);
        if (h->pin_aux) cudaFreeHost(h->pin_aux);
        h->pin_off = nullptr;
        h->pin_aux = nullptr;
        h->pin_slots = 0;
        // host-mapped: written by chunk_offsets_kernel, read by the host after an event (no copy-engine transfer)
        CK(cudaHostAlloc((void**)&h->pin_off, sizeof(long long) * (cslots + 1), cudaHostAllocMapped));
        CK(cudaHostAlloc((void**)&h->pin_aux, sizeof(int32_t) * 2 * cslots, cudaHostAllocMapped));
        h->pin_slots = cslots;
    }
    long long* pin_off_dev = nullptr;
    int32_t* pin_aux_dev = nullptr;
    CK(cudaHostGetDevicePointer((void**)&pin_off_dev, h->pin_off, 0));
    CK(cudaHostGetDevicePointer((void**)&pin_aux_dev, h->pin_aux, 0));
    if ((rc = ensure(h, h->pack_off, sizeof(long long) * (cslots + 1), false))) return rc;
    if (compact) {
        if ((rc = ensure(h, h->cmp_aux, sizeof(int32_t) * cslots, false))) return rc;
        if ((rc = ensure(h, h->cmp_wsize, sizeof(int32_t) * cslots, false))) return rc;
        if ((rc = ensure(h, h->cmp_yw, sizeof(int32_t) * cslots, false))) return rc;
        if ((rc = ensure(h, h->cmp_codes, sizeof(unsigned) * cslots * l.cap, false))) return rc;
    }
    // compact records never exceed 2 n + n / 64 + 8 words (full-width ordinates, one explicit run), packed lines 2 n
    const size_t pack_bytes = compact ? sizeof(double) * cslots * (2 * (size_t)l.cap + l.cap / 64 + 8) : sizeof(double2) * cslots * l.cap;
    if ((rc = ensure(h, h->pack_xy, pack_bytes, false))) return rc;
    if (nchunks > 1)
        if ((rc = ensure(h, h->pack_xy2, pack_bytes, false))) return rc;

    CK(cudaEventRecord(h->ev[2], h->stream));
    CK(cudaMemsetAsync(B.isects, 0, sizeof(double2) * nstates * l.isec_cap, h->stream));
    CK(cudaMemsetAsync(B.removed, 0, sizeof(uint32_t) * nstates * l.rem_words, h->stream));
    CK(cudaMemsetAsync(B.lens, 0, sizeof(int32_t) * nstates * 4, h->stream));
    CK(cudaMemsetAsync(B.meta, 0, sizeof(dcegm_env_meta) * nstates, h->stream));
    CK(cudaMemsetAsync(B.vbound, 0, sizeof(double) * nstates * 2, h->stream));
    int launches = 0;
    DeviceBatch Bc = B;
    Bc.order = nullptr;
    // chunks are contiguous model ranges (their lines are delivered in batch order); inside a chunk the costliest
    // models (smallest taste-shock scale) start first, so that the tail of every chunk's launch is made of cheap ones
    if (!h->big && nchunks > 1 && (model == DCEGM_MODEL_F || model == DCEGM_MODEL_G)) {
        std::vector<int32_t> ord(nmodels);
        for (int i = 0; i < nmodels; i++) ord[i] = i;
        for (int c = 0; c < nchunks; c++)
            std::stable_sort(ord.begin() + cstart[c], ord.begin() + cstart[c] + ccount[c],
                             [&](int32_t a, int32_t b) { return params[a].lambda < params[b].lambda; });
        if ((rc = ensure(h, h->order, sizeof(int32_t) * nmodels, false))) return rc;
        CK(cudaMemcpyAsync(h->order.p, ord.data(), sizeof(int32_t) * nmodels, cudaMemcpyHostToDevice, h->stream));
        CK(cudaStreamSynchronize(h->stream));  // ord goes out of scope
        Bc.order = (const int32_t*)h->order.p;
        h->batch.order = Bc.order;  // still a permutation of the batch: later whole-batch runs may use it
    }
    // ONE launch for the whole batch when the driver offers stream memory operations: finished models count up their
    // chunk's counter and the packing stream waits for the counter instead of for a per-chunk kernel — no drain of the
    // GPU at every chunk boundary.  DCEGM_ONE_LAUNCH=0 keeps one launch per chunk.
    bool one_launch = want_one && Bc.order != nullptr;
    if (one_launch) {
        std::vector<uint16_t> mc(nmodels);
        for (int c = 0; c < nchunks; c++)
            for (int i = 0; i < ccount[c]; i++) mc[cstart[c] + i] = (uint16_t)c;  // chunks are model ranges
        if ((rc = ensure(h, h->model_chunk, sizeof(uint16_t) * nmodels, false))) return rc;
        if ((rc = ensure(h, h->chunk_done, sizeof(int32_t) * nchunks, false))) return rc;
        CK(cudaMemcpyAsync(h->model_chunk.p, mc.data(), sizeof(uint16_t) * nmodels, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemsetAsync(h->chunk_done.p, 0, sizeof(int32_t) * nchunks, h->stream));
        CK(cudaStreamSynchronize(h->stream));  // mc goes out of scope; the counters are zero before anybody waits
        if (h->wait_value_state == 0) {  // first use: does the wait work on this device / driver?
            const CUresult r = wait_value((CUstream)h->s_pack, (CUdeviceptr)h->chunk_done.p, 0u, CU_STREAM_WAIT_VALUE_GEQ);
            h->wait_value_state = (r == CUDA_SUCCESS && cudaStreamSynchronize(h->s_pack) == cudaSuccess) ? 1 : -1;
            cudaGetLastError();
            one_launch = h->wait_value_state == 1;
        }
    }
    if (one_launch) {
        Bc.chunk_done = (int32_t*)h->chunk_done.p;
        Bc.model_chunk = (const uint16_t*)h->model_chunk.p;
        cudaError_t e = launch_range(h, Bc, 0, nmodels, h->stream);
        if (e != cudaSuccess) return cuda_fail(h, e, "kernel launch");
        launches += h->launches_last;
    }
    // one launch per chunk: the fine-grid solvers issue hundreds of launches per chunk, and more than ~1000 pending
    // launches make the host block in cudaLaunchKernel (it then sees a finished chunk's offsets late): only two chunks
    // are in the launch queue at a time there
    int launched = one_launch ? nchunks : 0;
    const int lookahead = h->big ? 2 : nchunks;
    auto launch_next = [&]() -> int {
        const int c = launched;
        cudaError_t e = launch_range(h, Bc, cstart[c], ccount[c], h->stream);
        if (e != cudaSuccess) return cuda_fail(h, e, "kernel launch");
        launches += h->launches_last;
        CK(cudaEventRecord(h->ev_pool[3 * c], h->stream));
        if (++launched == nchunks) CK(cudaEventRecord(h->ev[3], h->stream));
        return DCEGM_OK;
    };
    if (one_launch) CK(cudaEventRecord(h->ev[3], h->stream));
    while (launched < nchunks && launched < lookahead)
        if ((rc = launch_next())) return rc;

    int64_t base = 0;
    offsets[0] = 0;
    int err = DCEGM_OK;
    for (int c = 0; c < nchunks && err == DCEGM_OK; c++) {
        const int c0 = cstart[c], cnt = ccount[c];
        const size_t nsl = per * cnt;
        if (one_launch) {
            if (wait_value((CUstream)h->s_pack, (CUdeviceptr)((int32_t*)h->chunk_done.p + c), (cuuint32_t)cnt, CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
                return fail(h, DCEGM_E_CUDA, "cuStreamWaitValue32 failed");
        } else {
            CK(cudaStreamWaitEvent(h->s_pack, h->ev_pool[3 * c], 0));
        }
        if (c >= 1) CK(cudaStreamWaitEvent(h->s_pack, h->ev_pool[3 * (c - 1) + 1], 0));  // pack_off still read by pack c-1
        cudaError_t e;
        if (compact) {
            e = launch_compact_count(B, c0, cnt, compact_alias(), (int32_t*)h->cmp_aux.p, (int32_t*)h->cmp_yw.p, (int32_t*)h->cmp_wsize.p, (unsigned*)h->cmp_codes.p,
                                     pin_aux_dev, pin_aux_dev + cslots, h->s_pack);
            if (e != cudaSuccess) return cuda_fail(h, e, "compact count kernel");
            e = launch_chunk_offsets((const int32_t*)h->cmp_wsize.p, nsl, 0x7fffffff, (long long*)h->pack_off.p, pin_off_dev, h->s_pack);
        } else {
            e = launch_chunk_offsets(B.lens + per * c0, nsl, l.cap, (long long*)h->pack_off.p, pin_off_dev, h->s_pack);
        }
        if (e != cudaSuccess) return cuda_fail(h, e, "offsets kernel");
        CK(cudaStreamSynchronize(h->s_pack));
        if (trace) fprintf(stderr, "[dcegm] %8.3f ms chunk %d solved, offsets on host\n", trace_ms(), c);
        const long long run_ = h->pin_off[nsl];
        int64_t* go = offsets + per * c0;
        for (size_t s = 0; s <= nsl; s++) go[s] = base + h->pin_off[s];
        if (compact) {
            memcpy(aux_out + per * c0, h->pin_aux, sizeof(int32_t) * nsl);
            memcpy(lens_out + per * c0, h->pin_aux + cslots, sizeof(int32_t) * nsl);
            resolve_aliases(go, aux_out + per * c0, per, nsl);
        }
        if ((size_t)(base + run_) > cap_points) {
            err = fail(h, DCEGM_E_CAPACITY, "dcegm_solve_*_packed: destination too small");
            break;
        }
        double2* buf = (double2*)((c & 1) ? h->pack_xy2.p : h->pack_xy.p);
        if (c >= 2) CK(cudaStreamWaitEvent(h->s_pack, h->ev_pool[3 * (c - 2) + 2], 0));  // buffer free again
        if (compact)
            e = launch_compact_write(B, c0, cnt, (const int32_t*)h->cmp_aux.p, (const int32_t*)h->cmp_yw.p, (const unsigned*)h->cmp_codes.p,
                                     (const long long*)h->pack_off.p, (double*)buf, h->s_pack);
        else
            e = launch_pack_lines(B, c0, cnt, (const long long*)h->pack_off.p, buf, h->s_pack);
        if (e != cudaSuccess) return cuda_fail(h, e, "pack kernel");
        CK(cudaEventRecord(h->ev_pool[3 * c + 1], h->s_pack));
        CK(cudaStreamWaitEvent(h->s_copy, h->ev_pool[3 * c + 1], 0));
        const size_t unit = compact ? sizeof(double) : sizeof(double2);
        CK(cudaMemcpyAsync((char*)xy + unit * (size_t)base, buf, unit * (size_t)run_, cudaMemcpyDeviceToHost, h->s_copy));
        CK(cudaEventRecord(h->ev_pool[3 * c + 2], h->s_copy));
        base += run_;
        if (launched < nchunks && (rc = launch_next())) return rc;
    }
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->s_pack));
    if (trace) fprintf(stderr, "[dcegm] %8.3f ms all packed (%d chunks of %d)\n", trace_ms(), nchunks, chunk);
    CK(cudaStreamSynchronize(h->s_copy));
    if (trace) fprintf(stderr, "[dcegm] %8.3f ms all copied: %.1f MB\n", trace_ms(), base * (compact ? 8.0 : 16.0) / 1e6);
    CK(cudaEventElapsedTime(&h->timing.kernel_ms, h->ev[2], h->ev[3]));
    h->timing.n_launches = launches + (compact ? 3 : 2) * nchunks;
    h->solved = true;
    *npoints = base;
    return err;
}

int need_solution(dcegm_ctx* h) {
    if (!h) return DCEGM_E_INVALID;
    if (!h->solved) return fail(h, DCEGM_E_NOSOLUTION, "no resident solution: call dcegm_solve_* first");
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaSetDevice");
    return DCEGM_OK;
}

int check_range(dcegm_ctx* h, int first, int count) {
    if (first < 0 || count < 0 || first + count > h->batch.lay.nmodels) return fail(h, DCEGM_E_INVALID, "model range out of bounds");
    return DCEGM_OK;
}

template <typename T>
int fetch_per_model(dcegm_ctx* h, const void* base, size_t per_model, int first, int count, T* dst) {
    int rc = need_solution(h);
    if (rc) return rc;
    if ((rc = check_range(h, first, count))) return rc;
    if (!dst) return fail(h, DCEGM_E_INVALID, "destination is null");
    if (!base) return fail(h, DCEGM_E_NOSOLUTION, "this buffer was not produced by the last solve");
    CK(cudaMemcpyAsync(dst, (const char*)base + per_model * sizeof(T) * first, per_model * sizeof(T) * count,
                       cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

}  // namespace

extern "C" {

int dcegm_abi_version(void) { return DCEGM_ABI_VERSION; }

void dcegm_param_defaults(dcegm_param* p) {  // src/param.json
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->gamma = 1.95;
    p->beta = 0.95;
    p->sigma = 0.35;
    p->lambda = 2e-6;
    p->R = 1.0526315789473684;
    p->a_high = 50.0;
    p->a_low = 0.0;
    p->a_lowT = 0.0;
    p->cfloor = 0.001;
    p->cfloor_plot = 0.01;
    p->alpha = 0.35;
    p->alphaT = 0.0;
    p->inc0 = 0.75;
    p->inc1 = 0.04;
    p->inc2 = 0.0002;
    p->rho = 0.95;
    p->pension = 0.0;
    p->delta = 0.0;
    p->maxinterest = 200.0;
    p->bbar = 7.7;
    p->nu = 0.0;
    p->initw0 = 5.0;
    p->initw1 = 20.0;
    p->na = 100;
    p->ny = 5;
    p->nT = 25;
    p->nD = 2;
    p->k = 100;
    p->retage = 26;
    p->nsims = 100;
    p->quad_group = 1;
    p->rseed = 25032020;
}

int dcegm_create(int device_ordinal, dcegm_handle* out) {
    if (!out) return DCEGM_E_INVALID;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return DCEGM_E_NODEVICE;  // no CPU fallback
    if (device_ordinal < 0 || device_ordinal >= ndev) return DCEGM_E_INVALID;
    dcegm_ctx* h = new dcegm_ctx();
    h->device = device_ordinal;
    if (cudaSetDevice(device_ordinal) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete h;
        return DCEGM_E_CUDA;
    }
    for (auto& ev : h->ev) cudaEventCreate(&ev);
    cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device_ordinal);
    *out = h;
    return DCEGM_OK;
}

int dcegm_destroy(dcegm_handle h) {
    if (!h) return DCEGM_E_INVALID;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    DevBuf* bufs[] = {&h->models, &h->grids, &h->off, &h->wgt, &h->lines, &h->lens, &h->vbound, &h->meta, &h->isects,
                      &h->removed, &h->dirty, &h->status, &h->summaries, &h->pack_off, &h->pack_xy, &h->pack_xy2, &h->big_scratch, &h->cmp_aux, &h->cmp_wsize, &h->cmp_codes, &h->cmp_yw, &h->chunk_done, &h->model_chunk, &h->order,
                      &h->simm, &h->yvec, &h->mom, &h->rhist, &h->sstat, &h->pan_d, &h->pan_i, &h->pan_ret,
                      &h->sh_w, &h->sh_d, &h->sh_r, &h->sh_w0, &h->sh_y0};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (auto& ev : h->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : h->ev_pool) cudaEventDestroy(ev);
    if (h->pin_off) cudaFreeHost(h->pin_off);
    if (h->pin_aux) cudaFreeHost(h->pin_aux);
    for (auto& r : h->pin_tab)
        if (r.p) cudaHostUnregister(const_cast<void*>(r.p));
    if (h->s_pack_lo) cudaStreamDestroy(h->s_pack_lo);
    if (h->s_pack_hi) cudaStreamDestroy(h->s_pack_hi);
    if (h->s_copy) cudaStreamDestroy(h->s_copy);
    cudaStreamDestroy(h->stream);
    delete h;
    return DCEGM_OK;
}

const char* dcegm_last_error(dcegm_handle h) { return h ? h->err.c_str() : "null handle"; }

// Pinned host memory on the NUMA node of the current CUDA device: on a multi-GPU box every rank's device-to-host
// stream then lands in the memory behind its own root complex instead of wherever the process happened to start.
// Linux memory policy via raw syscalls (no libnuma in the image); failures are ignored (default placement).
// DCEGM_NUMA=0 switches it off.
int dcegm_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return DCEGM_E_INVALID;
    int node = -1;
    const char* env = getenv("DCEGM_NUMA");
    if (!env || atoi(env) != 0) {
        int dev = 0;
        char bus[32] = {0};
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetPCIBusId(bus, sizeof bus, dev) == cudaSuccess) {
            for (char* c = bus; *c; c++) *c = (char)tolower(*c);
            const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
            if (FILE* f = fopen(path.c_str(), "r")) {
                if (fscanf(f, "%d", &node) != 1) node = -1;
                fclose(f);
            }
        }
    }
    bool policy_set = false;
#if defined(__linux__) && defined(SYS_set_mempolicy)
    if (node >= 0 && node < 1024) {
        unsigned long mask[16] = {0};
        mask[node / (8 * sizeof(unsigned long))] = 1UL << (node % (8 * sizeof(unsigned long)));
        policy_set = syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, mask, (unsigned long)(8 * sizeof mask)) == 0;
    }
#endif
    const cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocDefault);
#if defined(__linux__) && defined(SYS_set_mempolicy)
    if (policy_set) syscall(SYS_set_mempolicy, 0 /* MPOL_DEFAULT */, nullptr, 0UL);
#endif
    if (getenv("DCEGM_TRACE")) fprintf(stderr, "[dcegm] host_alloc %zu B: numa node %d, policy %s\n", bytes, node, policy_set ? "preferred" : "default");
    return e == cudaSuccess ? DCEGM_OK : DCEGM_E_CUDA;
}
int dcegm_host_free(void* ptr) { return cudaFreeHost(ptr) == cudaSuccess ? DCEGM_OK : DCEGM_E_CUDA; }

int dcegm_upload_f(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) { return upload(h, DCEGM_MODEL_F, p, n, flags); }
int dcegm_upload_g(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) { return upload(h, DCEGM_MODEL_G, p, n, flags); }
int dcegm_upload_minimal(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    return upload(h, DCEGM_MODEL_MINIMAL, p, n, flags);
}
int dcegm_run(dcegm_handle h) { return run(h); }

int dcegm_solve_f(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    int rc = upload(h, DCEGM_MODEL_F, p, n, flags);
    return rc ? rc : run(h);
}
int dcegm_solve_g(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    int rc = upload(h, DCEGM_MODEL_G, p, n, flags);
    return rc ? rc : run(h);
}
int dcegm_solve_minimal_borrow(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    int rc = upload(h, DCEGM_MODEL_MINIMAL_BORROW, p, n, flags);
    return rc ? rc : run(h);
}
int dcegm_solve_minimal_bequest(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    int rc = upload(h, DCEGM_MODEL_MINIMAL_BEQUEST, p, n, flags);
    return rc ? rc : run(h);
}
int dcegm_solve_minimal(dcegm_handle h, const dcegm_param* p, int n, unsigned flags) {
    int rc = upload(h, DCEGM_MODEL_MINIMAL, p, n, flags);
    return rc ? rc : run(h);
}

int dcegm_solve_f_packed(dcegm_handle h, const dcegm_param* p, int n, unsigned flags, double* xy, size_t cap_points,
                         int64_t* offsets, int64_t* npoints) {
    return solve_packed(h, DCEGM_MODEL_F, p, n, flags, xy, cap_points, offsets, npoints);
}
int dcegm_solve_g_packed(dcegm_handle h, const dcegm_param* p, int n, unsigned flags, double* xy, size_t cap_points,
                         int64_t* offsets, int64_t* npoints) {
    return solve_packed(h, DCEGM_MODEL_G, p, n, flags, xy, cap_points, offsets, npoints);
}

int dcegm_get_layout(dcegm_handle h, dcegm_layout* out) {
    if (!h || !out) return DCEGM_E_INVALID;
    if (!h->uploaded) return fail(h, DCEGM_E_NOSOLUTION, "no batch uploaded");
    *out = h->batch.lay;
    return DCEGM_OK;
}

int dcegm_fetch_line(dcegm_handle h, int model, int kind, int id, int iy, int it, double* xy, int cap_points, int* len) {
    int rc = need_solution(h);
    if (rc) return rc;
    const dcegm_layout& l = h->batch.lay;
    if (model < 0 || model >= l.nmodels || id < 0 || id > 1 || iy < 0 || iy >= l.nstate || it < 0 || it >= l.nT || kind < 0 ||
        kind > 3 || !len)
        return fail(h, DCEGM_E_INVALID, "dcegm_fetch_line: index out of range");
    const bool dirty = kind >= 2;
    const size_t slot = ((((size_t)model * l.nT + it) * l.nstate + iy) * 2 + id) * 2 + (kind & 1);
    int n;
    const double2* src;
    if (dirty) {
        if (!h->batch.dirty) return fail(h, DCEGM_E_NOSOLUTION, "solve without DCEGM_KEEP_DIRTY");
        n = it == l.nT - 1 ? 0 : l.na;
        src = h->batch.dirty + slot * (size_t)l.na;
    } else {
        if (!h->batch.lines) return fail(h, DCEGM_E_NOSOLUTION, "solve with DCEGM_SUMMARY_ONLY keeps no lines");
        CK(cudaMemcpyAsync(&n, h->batch.lens + slot, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        src = h->batch.lines + slot * (size_t)l.cap;
    }
    *len = n;
    if (n > cap_points || !xy) return fail(h, DCEGM_E_CAPACITY, "dcegm_fetch_line: destination too small");
    if (n > 0) {
        CK(cudaMemcpyAsync(xy, src, sizeof(double2) * n, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return DCEGM_OK;
}

int dcegm_fetch_lines(dcegm_handle h, int first, int count, double* lines) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<double>(h, h->batch.lines, (size_t)l.nT * l.nstate * 4 * l.cap * 2, first, count, lines);
}
int dcegm_fetch_lens(dcegm_handle h, int first, int count, int32_t* lens) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<int32_t>(h, h->batch.lens, (size_t)l.nT * l.nstate * 4, first, count, lens);
}
int dcegm_fetch_vbound(dcegm_handle h, int first, int count, double* vbound) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<double>(h, h->batch.vbound, (size_t)l.nT * l.nstate * 2, first, count, vbound);
}
int dcegm_fetch_meta(dcegm_handle h, int first, int count, dcegm_env_meta* meta) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<dcegm_env_meta>(h, h->batch.meta, (size_t)l.nT * l.nstate, first, count, meta);
}
int dcegm_fetch_isects(dcegm_handle h, int first, int count, double* isects) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<double>(h, h->batch.isects, (size_t)l.nT * l.nstate * l.isec_cap * 2, first, count, isects);
}
int dcegm_fetch_removed(dcegm_handle h, int first, int count, uint32_t* removed) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<uint32_t>(h, h->batch.removed, (size_t)l.nT * l.nstate * l.rem_words, first, count, removed);
}
int dcegm_fetch_dirty(dcegm_handle h, int first, int count, double* dirty) {
    if (!h) return DCEGM_E_INVALID;
    const dcegm_layout& l = h->batch.lay;
    return fetch_per_model<double>(h, h->batch.dirty, (size_t)l.nT * l.nstate * 4 * l.na * 2, first, count, dirty);
}

int dcegm_fetch_packed(dcegm_handle h, int first, int count, double* xy, size_t cap_points, int64_t* offsets, int64_t* npoints) {
    int rc = need_solution(h);
    if (rc) return rc;
    if ((rc = check_range(h, first, count))) return rc;
    if (!offsets || !npoints) return fail(h, DCEGM_E_INVALID, "offsets/npoints is null");
    if (!h->batch.lines) return fail(h, DCEGM_E_NOSOLUTION, "solve with DCEGM_SUMMARY_ONLY keeps no lines");
    const dcegm_layout& l = h->batch.lay;
    const size_t per = (size_t)l.nT * l.nstate * 4, nsl = per * count;
    std::vector<int32_t> lens(nsl);
    CK(cudaMemcpyAsync(lens.data(), h->batch.lens + per * first, sizeof(int32_t) * nsl, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    offsets[0] = 0;
    for (size_t s = 0; s < nsl; s++) offsets[s + 1] = offsets[s] + (lens[s] < l.cap ? lens[s] : l.cap);
    *npoints = offsets[nsl];
    if (!xy) return DCEGM_OK;
    if ((size_t)*npoints > cap_points) return fail(h, DCEGM_E_CAPACITY, "dcegm_fetch_packed: destination too small");
    if ((rc = ensure(h, h->pack_off, sizeof(long long) * (nsl + 1), false))) return rc;
    if ((rc = ensure(h, h->pack_xy, sizeof(double2) * (size_t)*npoints, false))) return rc;
    static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
    CK(cudaMemcpyAsync(h->pack_off.p, offsets, sizeof(int64_t) * (nsl + 1), cudaMemcpyHostToDevice, h->stream));
    cudaError_t e = launch_pack_lines(h->batch, first, count, (const long long*)h->pack_off.p, (double2*)h->pack_xy.p, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "pack kernel");
    CK(cudaMemcpyAsync(xy, h->pack_xy.p, sizeof(double2) * (size_t)*npoints, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_solve_f_compact(dcegm_handle h, const dcegm_param* p, int n, unsigned flags, double* words, size_t cap_words,
                          int64_t* woffsets, int32_t* lens, int32_t* aux, int64_t* nwords) {
    return solve_packed(h, DCEGM_MODEL_F, p, n, flags, words, cap_words, woffsets, nwords, 1, lens, aux);
}
int dcegm_solve_g_compact(dcegm_handle h, const dcegm_param* p, int n, unsigned flags, double* words, size_t cap_words,
                          int64_t* woffsets, int32_t* lens, int32_t* aux, int64_t* nwords) {
    return solve_packed(h, DCEGM_MODEL_G, p, n, flags, words, cap_words, woffsets, nwords, 1, lens, aux);
}

int dcegm_fetch_compact(dcegm_handle h, int first, int count, double* words, size_t cap_words, int64_t* woffsets, int32_t* lens,
                        int32_t* aux, int64_t* nwords) {
    int rc = need_solution(h);
    if (rc) return rc;
    if ((rc = check_range(h, first, count))) return rc;
    if (!woffsets || !nwords || !lens || !aux) return fail(h, DCEGM_E_INVALID, "woffsets/lens/aux/nwords is null");
    if (!h->batch.lines) return fail(h, DCEGM_E_NOSOLUTION, "solve with DCEGM_SUMMARY_ONLY keeps no lines");
    const dcegm_layout& l = h->batch.lay;
    const size_t per = (size_t)l.nT * l.nstate * 4, nsl = per * count;
    woffsets[0] = 0;
    *nwords = 0;
    if (nsl == 0) return DCEGM_OK;
    if ((rc = ensure(h, h->cmp_aux, sizeof(int32_t) * nsl, false))) return rc;
    if ((rc = ensure(h, h->cmp_wsize, sizeof(int32_t) * nsl, false))) return rc;
    if ((rc = ensure(h, h->pack_off, sizeof(long long) * (nsl + 1), false))) return rc;
    if ((rc = ensure(h, h->cmp_codes, sizeof(unsigned) * nsl * l.cap, false))) return rc;
    if ((rc = ensure(h, h->cmp_yw, sizeof(int32_t) * nsl, false))) return rc;
    cudaError_t e = launch_compact_count(h->batch, first, count, compact_alias(), (int32_t*)h->cmp_aux.p, (int32_t*)h->cmp_yw.p, (int32_t*)h->cmp_wsize.p,
                                         (unsigned*)h->cmp_codes.p, nullptr, nullptr, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "compact count kernel");
    std::vector<int32_t> wsize(nsl);
    CK(cudaMemcpyAsync(wsize.data(), h->cmp_wsize.p, sizeof(int32_t) * nsl, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(aux, h->cmp_aux.p, sizeof(int32_t) * nsl, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(lens, h->batch.lens + per * first, sizeof(int32_t) * nsl, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (size_t s = 0; s < nsl; s++) {
        lens[s] = lens[s] < l.cap ? lens[s] : l.cap;
        woffsets[s + 1] = woffsets[s] + wsize[s];
    }
    *nwords = woffsets[nsl];
    if (words) {  // device copy of the prefix offsets (aliased slots are skipped by the write kernel)
        if ((size_t)*nwords > cap_words) return fail(h, DCEGM_E_CAPACITY, "dcegm_fetch_compact: destination too small");
        CK(cudaMemcpyAsync(h->pack_off.p, woffsets, sizeof(int64_t) * (nsl + 1), cudaMemcpyHostToDevice, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    resolve_aliases(woffsets, aux, per, nsl);
    if (!words) return DCEGM_OK;
    if ((rc = ensure(h, h->pack_xy, sizeof(double) * (size_t)*nwords, false))) return rc;
    e = launch_compact_write(h->batch, first, count, (const int32_t*)h->cmp_aux.p, (const int32_t*)h->cmp_yw.p, (const unsigned*)h->cmp_codes.p,
                             (const long long*)h->pack_off.p, (double*)h->pack_xy.p, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "compact write kernel");
    CK(cudaMemcpyAsync(words, h->pack_xy.p, sizeof(double) * (size_t)*nwords, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_expand_compact(dcegm_handle h, int first, int count, const double* words, const int64_t* woffsets, const int32_t* lens,
                         const int32_t* aux, double* xy, size_t cap_points, int64_t* offsets, int64_t* npoints, int nthreads) {
    if (!h) return DCEGM_E_INVALID;
    if (!h->uploaded) return fail(h, DCEGM_E_NOSOLUTION, "dcegm_expand_compact needs the batch of the last solve (savings grids)");
    if (!words || !woffsets || !lens || !aux || !offsets || !npoints) return fail(h, DCEGM_E_INVALID, "null argument");
    const dcegm_layout& l = h->batch.lay;
    if (first < 0 || count < 0 || first + count > l.nmodels) return fail(h, DCEGM_E_INVALID, "model range out of bounds");
    const int rc = expand_compact(h->tab, l.nT, l.nstate, first, count, words, woffsets, lens, aux, xy, cap_points, offsets, nthreads);
    *npoints = offsets[(size_t)l.nT * l.nstate * 4 * count];
    if (rc == -5) return fail(h, DCEGM_E_CAPACITY, "dcegm_expand_compact: destination too small");
    if (rc) return fail(h, DCEGM_E_INVALID, "dcegm_expand_compact: inconsistent compact record");
    return DCEGM_OK;
}

int dcegm_decode_compact_pair(const double* grid, int na, const double* rec_c, int n_c, int aux_c, const double* rec_v, int n_v,
                              int aux_v, double* xy_c, double* xy_v) {
    if (!grid || na < 1 || !xy_c || (n_c > 0 && !rec_c) || (n_v > 0 && (!rec_v || !xy_v))) return DCEGM_E_INVALID;
    return decode_compact_pair(grid, na, rec_c, n_c, aux_c, rec_v, n_v, aux_v, xy_c, xy_v) ? DCEGM_E_INVALID : DCEGM_OK;
}

int dcegm_get_grid(dcegm_handle h, int model, double* grid, int cap, int* na) {
    if (!h) return DCEGM_E_INVALID;
    if (!h->uploaded) return fail(h, DCEGM_E_NOSOLUTION, "no batch uploaded");
    if (model < 0 || model >= h->batch.lay.nmodels || !na) return fail(h, DCEGM_E_INVALID, "model out of range");
    const DevModel& p = h->tab.models[model];
    *na = p.na;
    if (!grid) return DCEGM_OK;
    if (cap < p.na) return fail(h, DCEGM_E_CAPACITY, "dcegm_get_grid: destination too small");
    memcpy(grid, h->tab.grids.data() + (size_t)p.grid * p.na, sizeof(double) * p.na);
    return DCEGM_OK;
}

// ---- diagnostics: guard bands behind every device buffer of the handle -----------------------------------------
int dcegm_debug_check_guards(dcegm_handle h, int* nbuffers, int* nbad) {
    if (!h || !nbuffers || !nbad) return DCEGM_E_INVALID;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    DevBuf* bufs[] = {&h->models, &h->grids, &h->off, &h->wgt, &h->lines, &h->lens, &h->vbound, &h->meta, &h->isects,
                      &h->removed, &h->dirty, &h->status, &h->summaries, &h->pack_off, &h->pack_xy, &h->pack_xy2, &h->big_scratch,
                      &h->cmp_aux, &h->cmp_wsize, &h->cmp_codes, &h->cmp_yw, &h->chunk_done, &h->model_chunk, &h->order,
                      &h->simm, &h->yvec, &h->mom, &h->rhist, &h->sstat, &h->pan_d, &h->pan_i, &h->pan_ret,
                      &h->sh_w, &h->sh_d, &h->sh_r, &h->sh_w0, &h->sh_y0};
    std::vector<unsigned char> g(kGuardBytes);
    *nbuffers = *nbad = 0;
    for (DevBuf* b : bufs) {
        if (!b->p) continue;
        ++*nbuffers;
        CK(cudaMemcpy(g.data(), (const char*)b->p + b->cap, kGuardBytes, cudaMemcpyDeviceToHost));
        for (unsigned char c : g)
            if (c != 0xA5) {
                ++*nbad;
                break;
            }
    }
    return DCEGM_OK;
}

// ---- diagnostics: the host-side model constructors (host_setup.cpp), callable without a device --------
int dcegm_debug_host_table(int which, int n, double p0, double p1, double p2, int iarg, double* out0, double* out1) {
    if (n < 1 || !out0) return DCEGM_E_INVALID;
    std::vector<double> a, b;
    switch (which) {
        case 0:  // range(p0, stop = p1, length = n)
            linrange(p0, p1, n, out0);
            return DCEGM_OK;
        case 1:  // quadpoints(n, 0, 1): nodes, weights
            if (!out1) return DCEGM_E_INVALID;
            gauss_legendre01(n, a, b);
            break;
        case 2:  // quantile(Normal(0, 1), p0)
            out0[0] = normal_quantile(p0);
            return DCEGM_OK;
        case 3:  // gausshermite(n): nodes, weights
            if (!out1) return DCEGM_E_INVALID;
            gauss_hermite(n, a, b);
            break;
        case 4:  // rouwenhorst(rho = p0, mu = p1, sigma = p2, n): support, transition matrix (row-major)
            if (!out1) return DCEGM_E_INVALID;
            rouwenhorst(p0, p1, p2, n, a, b);
            break;
        case 5:  // scaleGrid(p0, p1, n; logorder = iarg)
            scale_grid(p0, p1, n, iarg, out0);
            return DCEGM_OK;
        default:
            return DCEGM_E_INVALID;
    }
    memcpy(out0, a.data(), sizeof(double) * a.size());
    memcpy(out1, b.data(), sizeof(double) * b.size());
    return DCEGM_OK;
}

int dcegm_status(dcegm_handle h, int32_t* st) {
    if (!h) return DCEGM_E_INVALID;
    return fetch_per_model<int32_t>(h, h->batch.status, 1, 0, h->batch.lay.nmodels, st);
}
int dcegm_summaries(dcegm_handle h, dcegm_summary* dst) {
    if (!h) return DCEGM_E_INVALID;
    return fetch_per_model<dcegm_summary>(h, h->batch.summaries, 1, 0, h->batch.lay.nmodels, dst);
}
int dcegm_summaries_to_device(dcegm_handle h, void* dst_device, size_t bytes) {
    int rc = need_solution(h);
    if (rc) return rc;
    const size_t need = sizeof(dcegm_summary) * (size_t)h->batch.lay.nmodels;
    if (!dst_device || bytes < need) return fail(h, DCEGM_E_CAPACITY, "dcegm_summaries_to_device: destination too small");
    CK(cudaMemcpyAsync(dst_device, h->batch.summaries, need, cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

size_t dcegm_sweep_record_bytes(int nT_pad) {
    return sizeof(dcegm_summary) + 8 + sizeof(double) * (size_t)(nT_pad > 0 ? nT_pad : 0) * DCEGM_SIM_NMOM;
}

int dcegm_sweep_records_to_device(dcegm_handle h, void* dst_device, size_t bytes, int nT_pad) {
    int rc = need_solution(h);
    if (rc) return rc;
    const dcegm_layout& l = h->batch.lay;
    if (nT_pad < l.nT) return fail(h, DCEGM_E_INVALID, "dcegm_sweep_records_to_device: nT_pad < nT");
    const size_t rec = dcegm_sweep_record_bytes(nT_pad);
    if (!dst_device || bytes < rec * (size_t)l.nmodels) return fail(h, DCEGM_E_CAPACITY, "dcegm_sweep_records_to_device: destination too small");
    cudaError_t e = launch_sweep_records(h->batch.summaries, h->simulated ? (const double*)h->mom.p : nullptr,
                                         (const int32_t*)h->sstat.p, l.nmodels, l.nT, nT_pad, rec, (unsigned char*)dst_device, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "sweep record kernel");
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_simulate(dcegm_handle h, unsigned flags, const dcegm_sim_shocks* sh) {
    int rc = need_solution(h);
    if (rc) return rc;
    const DeviceBatch& B = h->batch;
    const dcegm_layout& l = B.lay;
    if (l.model != DCEGM_MODEL_F && l.model != DCEGM_MODEL_G) return fail(h, DCEGM_E_INVALID, "dcegm_simulate: the resident solution is not an FModel / GModel");
    if (!B.lines) return fail(h, DCEGM_E_NOSOLUTION, "dcegm_simulate needs the stored lines (solve without DCEGM_SUMMARY_ONLY)");
    // lines of one period that do not fit into shared memory are read in place (sim_kernel<.., GL = true>)
    if ((long long)sim_smem_bytes(l, true) > (long long)h->max_smem_optin) return fail(h, DCEGM_E_INVALID, "dcegm_simulate: nT * ny tables do not fit into shared memory");
    if (l.ny > 256 || l.nT > 65533) return fail(h, DCEGM_E_INVALID, "dcegm_simulate: ny <= 256 and nT <= 65533 (packed agent state)");
    const int n = l.nmodels, nT = l.nT;
    const bool panel = flags & DCEGM_SIM_PANEL;
    int nsims = sh ? sh->nsims : h->tab.sim[0].nsims;
    if (sh) {
        if (nsims <= 0 || !sh->wshocks || !sh->dshocks || !sh->w0shocks || (l.model == DCEGM_MODEL_G && !sh->ystate0))
            return fail(h, DCEGM_E_INVALID, "dcegm_simulate: incomplete shock arrays");
    } else {
        for (int m = 0; m < n; m++) {
            if (h->tab.sim[m].nsims <= 0) return fail(h, DCEGM_E_INVALID, "dcegm_simulate: nsims must be positive");
            if (panel && h->tab.sim[m].nsims != nsims) return fail(h, DCEGM_E_INVALID, "DCEGM_SIM_PANEL needs one nsims for all models");
        }
    }
    SimBatch Sb{};
    Sb.sim = (const SimModel*)h->simm.p;
    Sb.yvec = (const double*)h->yvec.p;
    if ((rc = ensure(h, h->mom, sizeof(double) * (size_t)n * nT * DCEGM_SIM_NMOM, false))) return rc;
    if ((rc = ensure(h, h->rhist, sizeof(int32_t) * (size_t)n * (nT + 1), false))) return rc;
    if ((rc = ensure(h, h->sstat, sizeof(int32_t) * (size_t)n, false))) return rc;
    Sb.moments = (double*)h->mom.p;
    Sb.ret_hist = (int32_t*)h->rhist.p;
    Sb.sim_status = (int32_t*)h->sstat.p;
    const size_t npan = (size_t)nsims * nT;
    if (panel) {
        if ((rc = ensure(h, h->pan_d, sizeof(double) * 6 * npan * n, false))) return rc;
        if ((rc = ensure(h, h->pan_i, sizeof(int32_t) * 2 * npan * n, false))) return rc;
        if ((rc = ensure(h, h->pan_ret, sizeof(int32_t) * (size_t)nsims * n, false))) return rc;
        Sb.pan_d = (double*)h->pan_d.p;
        Sb.pan_i = (int32_t*)h->pan_i.p;
        Sb.pan_ret = (int32_t*)h->pan_ret.p;
    }
    if (sh) {
        Sb.given = 1;
        Sb.nsims_given = nsims;
        auto up = [&](DevBuf& b, const void* src, size_t bytes, const void** dst) -> int {
            *dst = nullptr;
            if (!src) return DCEGM_OK;
            int r = ensure(h, b, bytes, false);
            if (r) return r;
            if (cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, h->stream) != cudaSuccess) return fail(h, DCEGM_E_CUDA, "shock upload");
            *dst = b.p;
            return DCEGM_OK;
        };
        if ((rc = up(h->sh_w, sh->wshocks, sizeof(double) * npan, (const void**)&Sb.wsh))) return rc;
        if ((rc = up(h->sh_d, sh->dshocks, sizeof(double) * npan, (const void**)&Sb.dsh))) return rc;
        if ((rc = up(h->sh_r, sh->rshocks, sizeof(double) * npan, (const void**)&Sb.rsh))) return rc;
        if ((rc = up(h->sh_w0, sh->w0shocks, sizeof(double) * nsims, (const void**)&Sb.w0sh))) return rc;
        if ((rc = up(h->sh_y0, sh->ystate0, sizeof(int32_t) * nsims, (const void**)&Sb.y0))) return rc;
    }
    CK(cudaEventRecord(h->ev[0], h->stream));
    cudaError_t e = launch_sim(B, Sb, 0, n, h->max_smem_optin, h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "simulation kernel launch");
    CK(cudaEventRecord(h->ev[1], h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaEventElapsedTime(&h->sim_ms, h->ev[0], h->ev[1]));
    h->simulated = true;
    h->sim_nsims = panel ? nsims : 0;
    return DCEGM_OK;
}

int dcegm_fetch_sim_moments(dcegm_handle h, int first, int count, double* moments, int32_t* ret_hist, int32_t* sim_status) {
    int rc = need_solution(h);
    if (rc) return rc;
    if (!h->simulated) return fail(h, DCEGM_E_NOSOLUTION, "no simulation: call dcegm_simulate first");
    if ((rc = check_range(h, first, count))) return rc;
    const int nT = h->batch.lay.nT;
    if (moments)
        CK(cudaMemcpyAsync(moments, (const double*)h->mom.p + (size_t)first * nT * DCEGM_SIM_NMOM,
                           sizeof(double) * (size_t)count * nT * DCEGM_SIM_NMOM, cudaMemcpyDeviceToHost, h->stream));
    if (ret_hist)
        CK(cudaMemcpyAsync(ret_hist, (const int32_t*)h->rhist.p + (size_t)first * (nT + 1), sizeof(int32_t) * (size_t)count * (nT + 1),
                           cudaMemcpyDeviceToHost, h->stream));
    if (sim_status)
        CK(cudaMemcpyAsync(sim_status, (const int32_t*)h->sstat.p + first, sizeof(int32_t) * (size_t)count, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_sim_moments_to_device(dcegm_handle h, void* dst_device, size_t bytes) {
    int rc = need_solution(h);
    if (rc) return rc;
    if (!h->simulated) return fail(h, DCEGM_E_NOSOLUTION, "no simulation: call dcegm_simulate first");
    const size_t need = sizeof(double) * (size_t)h->batch.lay.nmodels * h->batch.lay.nT * DCEGM_SIM_NMOM;
    if (!dst_device || bytes < need) return fail(h, DCEGM_E_CAPACITY, "dcegm_sim_moments_to_device: destination too small");
    CK(cudaMemcpyAsync(dst_device, h->mom.p, need, cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_fetch_sim_panel(dcegm_handle h, int model, double* w0, double* w1, double* cons, double* util, double* inc,
                          double* prob_work, int32_t* ystate, int32_t* worker, int32_t* ret_age) {
    int rc = need_solution(h);
    if (rc) return rc;
    if (!h->simulated || h->sim_nsims == 0) return fail(h, DCEGM_E_NOSOLUTION, "no panel: call dcegm_simulate with DCEGM_SIM_PANEL");
    if (model < 0 || model >= h->batch.lay.nmodels) return fail(h, DCEGM_E_INVALID, "model out of range");
    const size_t npan = (size_t)h->sim_nsims * h->batch.lay.nT;
    double* dst[6] = {w0, w1, cons, util, inc, prob_work};
    for (int k = 0; k < 6; k++)
        if (dst[k])
            CK(cudaMemcpyAsync(dst[k], (const double*)h->pan_d.p + ((size_t)model * 6 + k) * npan, sizeof(double) * npan,
                               cudaMemcpyDeviceToHost, h->stream));
    int32_t* di[2] = {ystate, worker};
    for (int k = 0; k < 2; k++)
        if (di[k])
            CK(cudaMemcpyAsync(di[k], (const int32_t*)h->pan_i.p + ((size_t)model * 2 + k) * npan, sizeof(int32_t) * npan,
                               cudaMemcpyDeviceToHost, h->stream));
    if (ret_age)
        CK(cudaMemcpyAsync(ret_age, (const int32_t*)h->pan_ret.p + (size_t)model * h->sim_nsims, sizeof(int32_t) * h->sim_nsims,
                           cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return DCEGM_OK;
}

int dcegm_last_timing(dcegm_handle h, dcegm_timing* out) {
    if (!h || !out) return DCEGM_E_INVALID;
    *out = h->timing;
    return DCEGM_OK;
}

}  // extern "C"
