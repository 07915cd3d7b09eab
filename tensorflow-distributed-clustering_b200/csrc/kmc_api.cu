// C ABI of libkmc_b200 (see include/kmc.h): context, workspace layout, host<->device staging and
// the per-iteration orchestration of the kernels.  No CPU fallback anywhere: every entry point
// launches CUDA kernels on an sm_100 device or returns an error.
#include "kmc_common.cuh"
#include "kmc_device.cuh"

#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace kmc {

static thread_local std::string g_last_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

static int64_t align256(int64_t v) { return (v + 255) & ~(int64_t)255; }

static void layout_workspace(kmc_ctx* c) {
    int64_t off = 0;
    auto take = [&](int64_t bytes) { int64_t o = off; off = align256(off + bytes); return o; };
    c->off_stats = take(sizeof(DeviceStats));
    c->off_centroids = take(c->K * c->D * 4);
    c->off_cprev = take(c->K * c->D * 4);
    c->off_shift_slots = take(kUpdateMaxBlocks * 8);
    c->off_partials = take((c->K * c->D + c->K + 2) * 8);
    if (c->path == KMC_PATH_SMALL)     // per-CTA partial values of the small-D kernel: [K*(D+1)+1][CTAs]
        c->off_small_slots = take((c->K * (c->D + 1) + 1) * small_max_ctas(c->sm_count) * 8);
    if (c->path != KMC_PATH_SMALL) {
        // counting-sort scratch, one block per row chunk: hist | cursor | offsets | segoff
        c->sort_cursor = align256(c->K * 4);
        c->sort_offsets = c->sort_cursor + align256(c->K * 4);
        c->sort_segoff = c->sort_offsets + align256((c->K + 1) * 4);
        c->sort_stride = c->sort_segoff + align256((c->K + 1) * 4);
        c->off_sort = take(c->sort_stride * (kMaxChunks + 1));    // last slot: the incremental pass
        c->off_labels = take(c->n * 4);
        c->off_labels2 = take(c->n * 4);
        c->off_perm = take(c->n * 4);
        c->chg_cap = c->n / 4;
        c->off_changed = take((c->chg_cap > 0 ? c->chg_cap : 1) * 4);
        c->off_running = take((c->K * c->D + c->K + 2) * 8);
    }
    if (c->path == KMC_PATH_TC) {
        c->off_xnorm = take(c->n * 8);          // per row: |X'| and |X' - X'_hi|, both rounded up
        c->off_ambmask = take((c->n / 32 + 2) * 4);   // one bit per row: not certified by the tensor kernel
        c->off_cnorm = take(c->K_pad * 4);
        c->off_cbias = take(c->K_pad * kBiasCols * 2);   // S*|c_k|^2 as fp16 pieces: one extra MMA k-step adds it
        c->off_cplanes = take(2 * c->K_pad * c->D_pad * 2);
        c->off_cplane16 = take(c->K_pad * c->D_pad * 2);     // fp16-accumulator tier: hi plane at its own scale
        c->off_cbias16 = take(c->K_pad * kBiasCols * 2);
        c->off_ccdist = (c->K <= 8192) ? take(c->K * c->K * 4) : -1;   // pruning table of the refine pass
        if (c->K <= 8192) c->off_overflow = take(c->n * 4);            // rows whose refine walk hit its cap
        c->off_nbr = (c->K <= 8192) ? take(c->K * c->K * 4) : -1;      // ... sorted nearest-first per centroid
        c->off_xext = take(c->n * 16 * 2);                 // per row: constants + fp16 pieces of |X'|^2 (fp16-accumulator tier)
        c->off_xhi = take(c->n * c->D_pad * 2);
        c->off_xlo = take(c->n * c->D_pad * 2);
    }
    c->need_bytes = off;
}

static int check_ready(kmc_ctx* c, bool need_centroids) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    if (!c->ws) KMC_FAIL(KMC_E_STATE, "workspace not bound (kmc_bind_workspace)");
    if (!c->points && c->n > 0) KMC_FAIL(KMC_E_STATE, "points not bound (kmc_bind_points)");
    if (need_centroids && !c->have_centroids) KMC_FAIL(KMC_E_STATE, "centroids not set (kmc_set_centroids)");
    if (c->streaming)
        KMC_FAIL(KMC_E_STATE, "streamed load incomplete: %lld of %lld rows delivered (kmc_accumulate_rows)",
                 (long long)c->stream_rows, (long long)c->n);
    return 0;
}

// Tensor path: (re)build the split centroid planes if the centroids changed outside kmc_update.
static int ensure_planes(kmc_ctx* c, cudaStream_t s) {
    if (c->path != KMC_PATH_TC || c->planes_ready) return 0;
    if (c->nbr_pending) {                        // a side-stream build for older centroids may still be writing
        KMC_CUDA(cudaStreamWaitEvent(s, c->nbr_ready, 0));
        c->nbr_pending = false;
    }
    KMC_TRY(launch_centroid_absmax(c, s));
    KMC_TRY(launch_prep_centroids(c, s));
    if (c->off_ccdist >= 0) KMC_TRY(launch_ccdist(c, s));
    c->planes_ready = true;
    return 0;
}

// exact re-decision of the rows the tensor kernel left in `worklist` (their number is in *n_work)
static int refine_worklist(kmc_ctx* c, int32_t* labels, const int32_t* worklist, const unsigned long long* n_work,
                           cudaStream_t s) {
    if (c->off_ccdist >= 0 && c->nbr_pending && s != c->side_stream) {
        KMC_CUDA(cudaStreamWaitEvent(s, c->nbr_ready, 0));     // neighbour lists built on the side stream (kmc_update)
        c->nbr_pending = false;
    }
    if (c->off_ccdist >= 0) {
        // worklist = perm + r0 of a row chunk: the overflow list and its counter of the same chunk
        const int64_t r0 = worklist - c->perm();
        const int ch = (int)(n_work - c->stats()->n_amb);
        return launch_refine(c, labels, worklist, n_work, c->at<int32_t>(c->off_overflow) + r0, &c->stats()->n_over[ch], s);
    }
    return launch_simt_assign(c, worklist, n_work, 0, labels, s);
}

// One tensor pass over rows [r0, r1) of row chunk `ch`.  Adaptive precision (KMC_PARAM_TC_TERMS = 0) is decided on
// the device, pass by pass, with no host round trip: the 1-term kernel runs unless the chunk is in a "skip" period,
// the gate kernel looks at what it left uncertified (tc_gate_kernel), and the 3-term kernel launched right behind
// redoes the chunk if the gate raised its flag.  All of them are enqueued every time; the ones not needed return at
// once.
static int tc_pass(kmc_ctx* c, int32_t* labels, int64_t r0, int64_t r1, int ch, cudaStream_t s) {
    unsigned long long* cnt = &c->stats()->n_amb[ch];
    const bool adaptive = c->tc_terms == 0 && c->off_nbr >= 0;
    // adaptive: the cheapest tier first -- fp16 accumulators where a K sweep has at least two accumulator tiles (with
    // one tile per row tile, config 3, its heavier end-of-tile work cancels the gain: 0.832 against 0.836 ms per
    // iteration), unless the device has switched that tier off for the chunk (DeviceStats::acc16_off), in which case
    // the fp32-accumulator 1-term kernel enqueued right behind it runs instead
    const bool acc16_policy = adaptive && c->tc_acc16 && c->K_pad >= 2 * kTcBlockN;
    const unsigned int* skip = adaptive ? &c->stats()->tc_skip[ch] : nullptr;
    if (acc16_policy) {
        const unsigned int* off = &c->stats()->acc16_off[ch];
        KMC_TRY(launch_tc_assign(c, labels, c->perm() + r0, cnt, nullptr, r0, r1, 16, skip, 0u, s, off, 0u));
        KMC_TRY(launch_tc_assign(c, labels, c->perm() + r0, cnt, nullptr, r0, r1, 1, skip, 0u, s, off, 1u));
    } else {
        const int first_terms = c->tc_terms ? c->tc_terms : (adaptive ? 1 : 3);
        KMC_TRY(launch_tc_assign(c, labels, c->perm() + r0, cnt, nullptr, r0, r1, first_terms, skip, 0u, s));
    }
    KMC_TRY(launch_compact_worklist(c, c->perm() + r0, cnt, r0, r1, skip, 0u, s));
    if (adaptive) {
        unsigned int* flag = &c->stats()->tc3_flag[ch];
        KMC_TRY(launch_tc_gate(c, ch, r1 - r0, acc16_policy, s));
        KMC_TRY(launch_tc_assign(c, labels, c->perm() + r0, cnt, nullptr, r0, r1, 3, flag, 1u, s));
        KMC_TRY(launch_compact_worklist(c, c->perm() + r0, cnt, r0, r1, flag, 1u, s));
    }
    return 0;
}

static int prof_mark(kmc_ctx* c, int slot, int which, cudaStream_t s) {
    if (!c->profile) return 0;
    cudaEvent_t& e = c->prof_ev[slot % kmc_ctx::kProfRing][which];
    if (!e) KMC_CUDA(cudaEventCreate(&e));
    KMC_CUDA(cudaEventRecord(e, s));
    return 0;
}

// labels of every row; `slot` >= 0 brackets the assign kernel and the refine with profile events
static int assign_into(kmc_ctx* c, int32_t* labels, cudaStream_t s, int slot = -1) {
    switch (c->path) {
        case KMC_PATH_SMALL:
            return launch_small(c, false, labels, c->centroids(), s);
        case KMC_PATH_SIMT:
            KMC_TRY(launch_simt_assign(c, nullptr, nullptr, c->n, labels, s));
            if (slot >= 0) KMC_TRY(prof_mark(c, slot, 1, s));
            return 0;
        case KMC_PATH_TC:
            KMC_TRY(ensure_planes(c, s));
            if (slot >= 0) KMC_TRY(prof_mark(c, slot, 0, s));    // after plane prep, before the tensor kernel
            KMC_CUDA(cudaMemsetAsync(c->stats()->n_amb, 0, sizeof(unsigned long long) * 2 * kMaxChunks, s));
            KMC_TRY(tc_pass(c, labels, 0, c->n, 0, s));
            if (slot >= 0) KMC_TRY(prof_mark(c, slot, 1, s));
            return refine_worklist(c, labels, c->perm(), &c->stats()->n_amb[0], s);
    }
    KMC_FAIL(KMC_E_INVALID, "bad kernel path %d", c->path);
}

// fp64 -> fp32 of a pageable host array straight into a pinned buffer, on several host threads: the rounding is the
// device's ((float)double, to nearest even), the PCIe copy moves half the bytes and needs no device-side pass
static void parallel_convert_f64(float* dst, const double* src, size_t count) {
    const size_t kMin = (size_t)1 << 18;
    unsigned nt = std::thread::hardware_concurrency();
    nt = nt == 0 ? 1 : (nt > 16 ? 16 : nt);
    if (count < 2 * kMin) nt = 1;
    else if (count / nt < kMin) nt = (unsigned)(count / kMin);
    auto work = [=](size_t b, size_t e) { for (size_t i = b; i < e; ++i) dst[i] = (float)src[i]; };
    if (nt <= 1) { work(0, count); return; }
    std::vector<std::thread> th;
    const size_t per = (count / nt + 15) & ~(size_t)15;
    for (unsigned t = 0; t < nt; ++t) {
        const size_t b = (size_t)t * per;
        if (b >= count) break;
        th.emplace_back(work, b, b + per > count ? count : b + per);
    }
    for (auto& t : th) t.join();
}

static void parallel_memcpy(void* dst, const void* src, size_t bytes) {
    const size_t kMin = (size_t)2 << 20;
    if (bytes < 2 * kMin) { memcpy(dst, src, bytes); return; }
    unsigned nt = std::thread::hardware_concurrency();
    nt = nt == 0 ? 1 : (nt > 16 ? 16 : nt);
    if (bytes / nt < kMin) nt = (unsigned)(bytes / kMin);
    std::vector<std::thread> th;
    const size_t per = (bytes / nt + 63) & ~(size_t)63;
    for (unsigned t = 0; t < nt; ++t) {
        const size_t b = (size_t)t * per;
        if (b >= bytes) break;
        const size_t len = (b + per > bytes) ? bytes - b : per;
        th.emplace_back([=] { memcpy((char*)dst + b, (const char*)src + b, len); });
    }
    for (auto& t : th) t.join();
}

}  // namespace kmc

using namespace kmc;

// pinned double buffer (and, for conversions on the device, two device buffers) of the staging paths;
// sized on first use (H2D of pageable rows: 64 MB chunks for the host copy threads; labels D2H: 16 MB)
static const int64_t kStageChunk = (int64_t)64 << 20;    // bytes per buffer, row staging
static const int64_t kLabelChunk = (int64_t)16 << 20;    // bytes per buffer, label delivery

static int ensure_staging(kmc_ctx* c, int64_t bytes, bool need_dev) {
    if (c->stage_bytes < bytes) {
        for (int i = 0; i < 2; ++i) {
            if (c->pinned[i]) { KMC_CUDA(cudaFreeHost(c->pinned[i])); c->pinned[i] = nullptr; }
            if (c->dev_stage[i]) { KMC_CUDA(cudaFree(c->dev_stage[i])); c->dev_stage[i] = nullptr; }
            KMC_CUDA(cudaHostAlloc(&c->pinned[i], (size_t)bytes, cudaHostAllocDefault));
            if (!c->pinned_free[i]) KMC_CUDA(cudaEventCreateWithFlags(&c->pinned_free[i], cudaEventDisableTiming));
        }
        c->stage_bytes = bytes;
    }
    if (need_dev && !c->dev_stage[0])
        for (int i = 0; i < 2; ++i) KMC_CUDA(cudaMalloc(&c->dev_stage[i], (size_t)c->stage_bytes));
    return 0;
}

extern "C" {

int kmc_version(void) { return 2000; }

const char* kmc_last_error(void) { return g_last_error.c_str(); }

int kmc_create(kmc_ctx** out, int device, int64_t n_rows, int64_t D, int64_t K, uint32_t flags) {
    if (!out) KMC_FAIL(KMC_E_INVALID, "out is null");
    *out = nullptr;
    if (n_rows < 0 || D < 1 || K < 1) KMC_FAIL(KMC_E_INVALID, "need n_rows >= 0, D >= 1, K >= 1");
    if (n_rows >= ((int64_t)1 << 31)) KMC_FAIL(KMC_E_INVALID, "at most 2^31-1 rows per shard");
    if (K * D >= ((int64_t)1 << 31)) KMC_FAIL(KMC_E_INVALID, "K*D too large");
    KMC_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    KMC_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        KMC_FAIL(KMC_E_DEVICE, "device %d is sm_%d%d; this library contains sm_100a code only (no fallback)",
                 device, prop.major, prop.minor);
    kmc_ctx* c = new (std::nothrow) kmc_ctx();
    if (!c) KMC_FAIL(KMC_E_NOMEM, "host allocation failed");
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->n = n_rows; c->n_cap = n_rows; c->D = D; c->K = K; c->flags = flags;
    int path = (int)(flags & KMC_PATH_MASK);
    if (path == KMC_PATH_AUTO) {
        if (small_path_supported(K, D)) path = KMC_PATH_SMALL;
        else if (tc_path_supported(K, D) && K * D >= 4096) path = KMC_PATH_TC;
        else path = KMC_PATH_SIMT;
    } else if (path == KMC_PATH_SMALL && !small_path_supported(K, D)) {
        delete c;
        KMC_FAIL(KMC_E_UNSUPPORTED, "small path needs D <= 8 and K*(D+1) <= 384");
    } else if (path == KMC_PATH_TC && !tc_path_supported(K, D)) {
        delete c;
        KMC_FAIL(KMC_E_UNSUPPORTED, "tensor path needs 16 <= D <= 256 and K >= 32");
    } else if (path != KMC_PATH_SMALL && path != KMC_PATH_SIMT && path != KMC_PATH_TC) {
        delete c;
        KMC_FAIL(KMC_E_INVALID, "unknown kernel path flag 0x%x", path);
    }
    c->path = path;
    if (path == KMC_PATH_TC) {
        c->D_pad = round_up(D, kTcChunkK);
        if (c->D_pad == 3 * kTcChunkK) c->D_pad = 4 * kTcChunkK;
        c->K_pad = round_up(K, kTcBlockN);
        // error of one score relative to |X'_i|*|C'_k|: operand split (3 * 2^-22) plus one
        // truncation of the fp32 accumulator per MMA k-step, with 25% head-room
        const double steps = 3.0 * (double)(c->D_pad / 16);
        c->tc_gamma = 1.25 * (steps + 6.0) / 8388608.0;
    }
    layout_workspace(c);
    if (path == KMC_PATH_TC) {
        // one tensor term + exact refine (fewer MMAs, less power, cheaper epilogue) whenever the refine pass
        // has its sorted neighbour lists; the choice is revisited from the refine counts
        cudaError_t e1 = cudaSuccess, e2 = cudaSuccess;
        cudaError_t e3 = cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking);
        for (int i = 0; i < kMaxChunks && e3 == cudaSuccess; ++i)
            e3 = cudaEventCreateWithFlags(&c->chunk_done[i], cudaEventDisableTiming);
        if (e3 == cudaSuccess) e3 = cudaEventCreateWithFlags(&c->side_done, cudaEventDisableTiming);
        if (e3 == cudaSuccess) e3 = cudaEventCreateWithFlags(&c->side_go, cudaEventDisableTiming);
        if (e3 == cudaSuccess) e3 = cudaEventCreateWithFlags(&c->nbr_ready, cudaEventDisableTiming);
        if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
            kmc_destroy(c);
            KMC_FAIL((int)(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3)), "could not create streams / events");
        }
        c->n_chunks = 1;      // KMC_PARAM_CHUNKS > 1: the refine of row chunk c runs on the side stream under the tensor kernel of chunk c+1
    }
    *out = c;
    return 0;
}

void kmc_destroy(kmc_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (int i = 0; i < 2; ++i) {
        if (c->pinned[i]) cudaFreeHost(c->pinned[i]);
        if (c->pinned_free[i]) cudaEventDestroy(c->pinned_free[i]);
        if (c->dev_stage[i]) cudaFree(c->dev_stage[i]);
    }
    if (c->side_stream) { cudaStreamSynchronize(c->side_stream); cudaStreamDestroy(c->side_stream); }
    for (auto& e : c->chunk_done)
        if (e) cudaEventDestroy(e);
    if (c->side_done) cudaEventDestroy(c->side_done);
    if (c->side_go) cudaEventDestroy(c->side_go);
    if (c->nbr_ready) cudaEventDestroy(c->nbr_ready);
    for (auto& ring : c->prof_ev)
        for (auto& e : ring)
            if (e) cudaEventDestroy(e);
    delete c;
}

int64_t kmc_workspace_bytes(const kmc_ctx* c) { return c ? c->need_bytes : 0; }

int kmc_bind_workspace(kmc_ctx* c, void* dev_workspace, int64_t bytes) {
    if (!c || !dev_workspace) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (bytes < c->need_bytes)
        KMC_FAIL(KMC_E_NOMEM, "workspace of %lld bytes given, %lld needed", (long long)bytes, (long long)c->need_bytes);
    if (reinterpret_cast<uintptr_t>(dev_workspace) & 255) KMC_FAIL(KMC_E_INVALID, "workspace must be 256-byte aligned");
    KMC_CUDA(cudaSetDevice(c->device));
    c->ws = static_cast<char*>(dev_workspace);
    c->ws_bytes = bytes;
    KMC_CUDA(cudaMemset(c->stats(), 0, sizeof(DeviceStats)));
    c->have_centroids = false;
    c->planes_ready = false;
    c->tmaps_ready = false;
    c->delta_valid = false;
    c->lab_cur = 0;
    return 0;
}

int kmc_bind_points(kmc_ctx* c, const float* dev_points, void* stream) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    if (!c->ws) KMC_FAIL(KMC_E_STATE, "bind the workspace before the points");
    if (!dev_points && c->n > 0) KMC_FAIL(KMC_E_INVALID, "null points");
    if (reinterpret_cast<uintptr_t>(dev_points) & 3) KMC_FAIL(KMC_E_INVALID, "points must be 4-byte aligned");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    c->points = dev_points;
    c->planes_ready = false;
    c->streaming = false;
    c->delta_valid = false;
    if (c->path == KMC_PATH_TC) {
        KMC_CUDA(cudaMemsetAsync(&c->stats()->x_absmax_bits, 0, sizeof(unsigned int), s));
        KMC_TRY(launch_absmax(c, dev_points, c->n * c->D, &c->stats()->x_absmax_bits, s));
        KMC_TRY(launch_split_points(c, s));
        KMC_TRY(tc_setup_tensor_maps(c));
    }
    return 0;
}

int kmc_bind_points_streamed(kmc_ctx* c, const float* dev_points, void* stream) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    if (!c->ws) KMC_FAIL(KMC_E_STATE, "bind the workspace before the points");
    if (!dev_points && c->n > 0) KMC_FAIL(KMC_E_INVALID, "null points");
    if (reinterpret_cast<uintptr_t>(dev_points) & 3) KMC_FAIL(KMC_E_INVALID, "points must be 4-byte aligned");
    if (c->path != KMC_PATH_TC || c->n < kTcBlockM)
        KMC_FAIL(KMC_E_UNSUPPORTED, "streamed loading exists for the tensor path only; use kmc_bind_points");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    c->points = dev_points;
    c->planes_ready = false;
    c->streaming = true;
    c->stream_rows = 0;
    c->delta_valid = false;
    KMC_CUDA(cudaMemsetAsync(&c->stats()->x_absmax_bits, 0, sizeof(unsigned int), s));
    KMC_TRY(tc_setup_tensor_maps(c));
    return 0;
}

int kmc_accumulate_rows(kmc_ctx* c, int64_t r0, int64_t r1, void* stream) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    if (!c->streaming) KMC_FAIL(KMC_E_STATE, "kmc_accumulate_rows needs kmc_bind_points_streamed first");
    if (!c->have_centroids) KMC_FAIL(KMC_E_STATE, "centroids not set (kmc_set_centroids)");
    if (r0 != c->stream_rows || r1 <= r0 || r1 > c->n || (r1 < c->n && r1 % kTcBlockM))
        KMC_FAIL(KMC_E_INVALID, "rows [%lld, %lld): pieces are consecutive from row %lld and end at multiples of %d",
                 (long long)r0, (long long)r1, (long long)c->stream_rows, kTcBlockM);
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool first = r0 == 0;
    if (first) KMC_CUDA(cudaMemsetAsync(c->partials(), 0, (size_t)kmc_partials_len(c) * 8, s));
    // the first piece fixes the power-of-two scale of the fp16 planes (|X'| < 2^12 leaves 16x head-room in
    // fp16 for larger values in later pieces; KMC_STAT_X_SCALED_MAX tells the caller whether that was enough)
    KMC_TRY(launch_absmax(c, c->points + r0 * c->D, (r1 - r0) * c->D, &c->stats()->x_absmax_bits, s));
    KMC_TRY(launch_split_points_range(c, r0, r1, !first, s));
    KMC_TRY(ensure_planes(c, s));
    KMC_CUDA(cudaMemsetAsync(c->stats()->n_amb, 0, sizeof(unsigned long long) * 2 * kMaxChunks, s));
    KMC_TRY(tc_pass(c, c->labels(), r0, r1, 0, s));
    KMC_TRY(refine_worklist(c, c->labels(), c->perm() + r0, &c->stats()->n_amb[0], s));
    KMC_TRY(launch_sorted_accumulate(c, c->labels(), r0, r1, 0, nullptr, s));
    c->stream_rows = r1;
    if (r1 == c->n) {
        c->streaming = false;
        // same state as after kmc_accumulate: the sums of all pieces become the running state
        if (c->delta_enabled) {
            KMC_CUDA(cudaMemcpyAsync(c->running(), c->partials(), (size_t)kmc_partials_len(c) * 8, cudaMemcpyDeviceToDevice, s));
            c->delta_valid = true;
        }
        c->last_pass_host_full = true;
        c->swap_labels();
    }
    return 0;
}

int kmc_stage_points(kmc_ctx* c, float* dev_dst, const void* host_src, int src_dtype, int64_t row0,
                     int64_t n_rows, void* stream) {
    if (!c || (!dev_dst && n_rows > 0) || (!host_src && n_rows > 0)) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (src_dtype != KMC_DTYPE_F32 && src_dtype != KMC_DTYPE_F64) KMC_FAIL(KMC_E_INVALID, "bad src_dtype %d", src_dtype);
    if (n_rows <= 0) return 0;
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t esz = src_dtype == KMC_DTYPE_F64 ? 8 : 4;
    const char* src = static_cast<const char*>(host_src) + (size_t)row0 * c->D * esz;
    const int64_t count = n_rows * c->D;

    cudaPointerAttributes attr;
    bool src_pinned = false;
    if (cudaPointerGetAttributes(&attr, src) == cudaSuccess) src_pinned = (attr.type == cudaMemoryTypeHost);
    cudaGetLastError();   // a pageable pointer may leave a sticky-free error code behind

    if (src_pinned && src_dtype == KMC_DTYPE_F32) {
        KMC_CUDA(cudaMemcpyAsync(dev_dst, src, (size_t)count * 4, cudaMemcpyHostToDevice, s));
        KMC_CUDA(cudaStreamSynchronize(s));
        return 0;
    }
    const int64_t kChunk = kStageChunk;
    if (!src_pinned && src_dtype == KMC_DTYPE_F64) {
        // pageable fp64 rows (the reference's callers' arrays): converted to fp32 by the host threads that have to
        // touch every byte anyway, 64 MB of floats per pinned buffer
        KMC_TRY(ensure_staging(c, kStageChunk, false));
        const int64_t per = kChunk / 4;
        const double* src64 = reinterpret_cast<const double*>(src);
        int b = 0;
        for (int64_t e0 = 0; e0 < count; e0 += per, b ^= 1) {
            const int64_t ne = (count - e0 < per) ? count - e0 : per;
            KMC_CUDA(cudaEventSynchronize(c->pinned_free[b]));
            parallel_convert_f64(static_cast<float*>(c->pinned[b]), src64 + e0, (size_t)ne);
            KMC_CUDA(cudaMemcpyAsync(dev_dst + e0, c->pinned[b], (size_t)ne * 4, cudaMemcpyHostToDevice, s));
            KMC_CUDA(cudaEventRecord(c->pinned_free[b], s));
        }
        return 0;
    }
    KMC_TRY(ensure_staging(c, kStageChunk, src_dtype == KMC_DTYPE_F64));
    const int64_t elems_per_chunk = kChunk / (int64_t)esz;
    int b = 0;
    for (int64_t e0 = 0; e0 < count; e0 += elems_per_chunk, b ^= 1) {
        const int64_t ne = (count - e0 < elems_per_chunk) ? count - e0 : elems_per_chunk;
        KMC_CUDA(cudaEventSynchronize(c->pinned_free[b]));
        if (src_pinned) {
            // pinned fp64 source: DMA straight from the caller's buffer
            KMC_CUDA(cudaMemcpyAsync(c->dev_stage[b], src + (size_t)e0 * esz, (size_t)ne * esz, cudaMemcpyHostToDevice, s));
        } else {
            parallel_memcpy(c->pinned[b], src + (size_t)e0 * esz, (size_t)ne * esz);
            void* dst = (src_dtype == KMC_DTYPE_F64) ? c->dev_stage[b] : (void*)(dev_dst + e0);
            KMC_CUDA(cudaMemcpyAsync(dst, c->pinned[b], (size_t)ne * esz, cudaMemcpyHostToDevice, s));
        }
        if (src_dtype == KMC_DTYPE_F64) {
            KMC_TRY(launch_convert_f64(c, static_cast<const double*>(c->dev_stage[b]), dev_dst + e0, ne, s));
            c->launches++;
        }
        KMC_CUDA(cudaEventRecord(c->pinned_free[b], s));
    }
    if (src_pinned) KMC_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int kmc_set_centroids(kmc_ctx* c, const float* host_centroids, void* stream) {
    if (!c || !host_centroids) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (!c->ws) KMC_FAIL(KMC_E_STATE, "workspace not bound");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KMC_CUDA(cudaMemcpyAsync(c->centroids(), host_centroids, (size_t)(c->K * c->D) * 4, cudaMemcpyHostToDevice, s));
    KMC_CUDA(cudaStreamSynchronize(s));
    c->have_centroids = true;
    c->planes_ready = false;
    return 0;
}

int kmc_get_centroids(kmc_ctx* c, float* host_centroids, void* stream) {
    if (!c || !host_centroids) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (!c->ws || !c->have_centroids) KMC_FAIL(KMC_E_STATE, "no centroids yet");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KMC_CUDA(cudaMemcpyAsync(host_centroids, c->centroids(), (size_t)(c->K * c->D) * 4, cudaMemcpyDeviceToHost, s));
    KMC_CUDA(cudaStreamSynchronize(s));
    return 0;
}

float* kmc_centroids_ptr(kmc_ctx* c) { return (c && c->ws) ? c->centroids() : nullptr; }
double* kmc_partials_ptr(kmc_ctx* c) { return (c && c->ws) ? c->partials() : nullptr; }
int64_t kmc_partials_len(const kmc_ctx* c) { return c ? c->K * c->D + c->K + 2 : 0; }

int kmc_resize(kmc_ctx* c, int64_t n_rows) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    if (n_rows < 0 || n_rows > c->n_cap)
        KMC_FAIL(KMC_E_INVALID, "kmc_resize: %lld rows outside [0, %lld], the size the context was created for",
                 (long long)n_rows, (long long)c->n_cap);
    c->n = n_rows;
    c->points = nullptr;          // the next batch must be bound before use
    c->streaming = false;
    c->tmaps_ready = false;
    c->delta_valid = false;
    return 0;
}

static int accumulate_impl(kmc_ctx* c, void* stream, bool zero);

int kmc_accumulate(kmc_ctx* c, void* stream) { return accumulate_impl(c, stream, true); }

int kmc_accumulate_add(kmc_ctx* c, void* stream) { return accumulate_impl(c, stream, false); }

static int accumulate_impl(kmc_ctx* c, void* stream, bool zero) {
    KMC_TRY(check_ready(c, true));
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (zero) KMC_CUDA(cudaMemsetAsync(c->partials(), 0, (size_t)kmc_partials_len(c) * 8, s));
    if (c->n == 0) return 0;
    const int slot = c->prof_acc_n;
    KMC_TRY(prof_mark(c, slot, 0, s));
    if (c->path == KMC_PATH_SMALL) {
        KMC_TRY(launch_small(c, true, nullptr, c->centroids(), s));
        KMC_TRY(prof_mark(c, slot, 1, s));
        KMC_TRY(prof_mark(c, slot, 2, s));
    } else {
        if (c->path == KMC_PATH_TC && c->n_chunks > 1 && c->n >= 4 * kTcBlockM) {
            // chunked tensor path: the caller's stream runs the tensor kernel over one row chunk after the other;
            // the side stream follows one chunk behind with the exact refine of the finished chunk (gather
            // latency there, tensor pipe here)
            KMC_TRY(ensure_planes(c, s));
            KMC_TRY(prof_mark(c, slot, 0, s));
            KMC_CUDA(cudaMemsetAsync(c->stats()->n_amb, 0, sizeof(unsigned long long) * 2 * kMaxChunks, s));
            KMC_CUDA(cudaEventRecord(c->side_go, s));                 // centroid planes and neighbour lists ready
            KMC_CUDA(cudaStreamWaitEvent(c->side_stream, c->side_go, 0));
            const int64_t tiles = (c->n + kTcBlockM - 1) / kTcBlockM;
            const int64_t tiles_per_chunk = (tiles + c->n_chunks - 1) / c->n_chunks;
            for (int ch = 0; ch < c->n_chunks; ++ch) {
                const int64_t r0 = (int64_t)ch * tiles_per_chunk * kTcBlockM;
                const int64_t r1 = (r0 + tiles_per_chunk * kTcBlockM < c->n) ? r0 + tiles_per_chunk * kTcBlockM : c->n;
                if (r0 >= c->n) break;
                KMC_TRY(tc_pass(c, c->labels(), r0, r1, ch, s));
                KMC_CUDA(cudaEventRecord(c->chunk_done[ch], s));
                KMC_CUDA(cudaStreamWaitEvent(c->side_stream, c->chunk_done[ch], 0));
                KMC_TRY(refine_worklist(c, c->labels(), c->perm() + r0, &c->stats()->n_amb[ch], c->side_stream));
            }
            KMC_TRY(prof_mark(c, slot, 1, s));                        // end of the last tensor kernel
            KMC_CUDA(cudaEventRecord(c->side_done, c->side_stream));
            KMC_CUDA(cudaStreamWaitEvent(s, c->side_done, 0));        // phase "refine" = its exposed tail
        } else {
            KMC_TRY(assign_into(c, c->labels(), s, slot));
        }
        KMC_TRY(prof_mark(c, slot, 2, s));
        if (zero) {
            c->last_pass_host_full = !(c->delta_enabled && c->delta_valid);
            KMC_TRY(launch_accumulate_stage(c, s));
        } else {
            // a batch of an out-of-core pass: its sums join those of other batches, no running state to keep
            KMC_TRY(launch_sorted_accumulate(c, c->labels(), 0, c->n, 0, nullptr, s));
            c->delta_valid = false;
            c->last_pass_host_full = true;
            c->swap_labels();
        }
    }
    KMC_TRY(prof_mark(c, slot, 3, s));
    if (c->profile) c->prof_acc_n++;
    c->updated_since_acc = false;
    return 0;
}

int kmc_update(kmc_ctx* c, void* stream) {
    KMC_TRY(check_ready(c, true));
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int slot = c->prof_upd_n;
    KMC_TRY(prof_mark(c, slot, 4, s));
    KMC_TRY(launch_update(c, s));
    c->updated_since_acc = true;
    if (c->path == KMC_PATH_TC) {
        KMC_TRY(launch_prep_centroids(c, s));   // update_kernel refreshed c_absmax
        if (c->off_ccdist >= 0) {
            // centroid-centroid distances and sorted neighbour lists are needed by the refine only: they are
            // built on the side stream while the next tensor pass runs (refine_worklist waits for nbr_ready)
            KMC_CUDA(cudaEventRecord(c->side_go, s));
            KMC_CUDA(cudaStreamWaitEvent(c->side_stream, c->side_go, 0));
            KMC_TRY(launch_ccdist(c, c->side_stream));
            KMC_CUDA(cudaEventRecord(c->nbr_ready, c->side_stream));
            c->nbr_pending = true;
        }
        c->planes_ready = true;
    }
    KMC_TRY(prof_mark(c, slot, 5, s));
    if (c->profile) c->prof_upd_n++;
    return 0;
}

int kmc_read_stats(kmc_ctx* c, double* host_stats, void* stream) {
    if (!c || !host_stats) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (!c->ws) KMC_FAIL(KMC_E_STATE, "workspace not bound");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DeviceStats st;
    double inertia = 0.0;
    KMC_CUDA(cudaMemcpyAsync(&st, c->stats(), sizeof(st), cudaMemcpyDeviceToHost, s));
    KMC_CUDA(cudaMemcpyAsync(&inertia, c->partials() + c->K * c->D + c->K, 8, cudaMemcpyDeviceToHost, s));
    KMC_CUDA(cudaStreamSynchronize(s));
    if (st.watchdog) KMC_FAIL(KMC_E_DEVICE, "tensor kernel barrier wait timed out (code 0x%x)", st.watchdog);
    for (int i = 0; i < KMC_STAT_COUNT; ++i) host_stats[i] = 0.0;
    host_stats[KMC_STAT_INERTIA] = inertia;
    host_stats[KMC_STAT_SHIFT2] = st.shift2;
    unsigned long long amb = 0;
    for (int i = 0; i < kMaxChunks; ++i) amb += st.n_amb[i];
    host_stats[KMC_STAT_AMBIGUOUS] = (double)amb;
    host_stats[KMC_STAT_KERNEL_PATH] = (double)c->path;
    host_stats[KMC_STAT_LAUNCHES] = (double)c->launches;
    host_stats[KMC_STAT_X_SCALED_MAX] = (double)__builtin_bit_cast(float, st.x_absmax_bits) * (double)st.sigma_x;
    unsigned long long over = 0;
    for (int i = 0; i < kMaxChunks; ++i) over += st.n_over[i];
    host_stats[KMC_STAT_WALK_OVERFLOW] = (double)over;
    host_stats[KMC_STAT_CHANGED] = c->last_pass_host_full ? (double)c->n : (double)st.n_changed;
    host_stats[KMC_STAT_FULL_PASS] = (c->last_pass_host_full || st.acc_full) ? 1.0 : 0.0;
    return 0;
}

int kmc_iterate(kmc_ctx* c, int max_iters, double tol, int* n_iter_out, double* inertia_out,
                double* shift2_out, void* stream) {
    KMC_TRY(check_ready(c, true));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int it = 0;
    if (c->path == KMC_PATH_SMALL && c->n > 0 && max_iters > 0) {
        // small-D path: one launch per iteration (the last CTA of the fused kernel folds the per-CTA sums into the
        // partial buffer and updates the centroids)
        KMC_CUDA(cudaSetDevice(c->device));
        for (; it < max_iters;) {
            KMC_TRY(launch_small_into(c, true, nullptr, c->centroids(), s, c->partials(), true));
            ++it;
            if (tol > 0.0) {
                double shift2 = 0.0;
                KMC_CUDA(cudaMemcpyAsync(&shift2, &c->stats()->shift2, 8, cudaMemcpyDeviceToHost, s));
                KMC_CUDA(cudaStreamSynchronize(s));
                if (shift2 <= tol) break;
            }
        }
        c->updated_since_acc = true;
        max_iters = 0;    // fall through to the common epilogue
    }
    for (; it < max_iters;) {
        KMC_TRY(kmc_accumulate(c, stream));
        KMC_TRY(kmc_update(c, stream));
        ++it;
        if (tol > 0.0) {
            double shift2 = 0.0;
            KMC_CUDA(cudaMemcpyAsync(&shift2, &c->stats()->shift2, 8, cudaMemcpyDeviceToHost, s));
            KMC_CUDA(cudaStreamSynchronize(s));
            if (shift2 <= tol) break;
        }
    }
    double st[KMC_STAT_COUNT];
    KMC_TRY(kmc_read_stats(c, st, stream));
    if (n_iter_out) *n_iter_out = it;
    if (inertia_out) *inertia_out = st[KMC_STAT_INERTIA];
    if (shift2_out) *shift2_out = st[KMC_STAT_SHIFT2];
    return 0;
}

int kmc_assign(kmc_ctx* c, int32_t* dev_labels, void* stream) {
    KMC_TRY(check_ready(c, true));
    if (!dev_labels && c->n > 0) KMC_FAIL(KMC_E_INVALID, "null labels");
    if (c->n == 0) return 0;
    KMC_CUDA(cudaSetDevice(c->device));
    return assign_into(c, dev_labels, static_cast<cudaStream_t>(stream));
}

int kmc_last_labels(kmc_ctx* c, int32_t* dev_labels, void* stream) {
    KMC_TRY(check_ready(c, true));
    if (!dev_labels && c->n > 0) KMC_FAIL(KMC_E_INVALID, "null labels");
    if (c->n == 0) return 0;
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (c->path != KMC_PATH_SMALL) {
        KMC_CUDA(cudaMemcpyAsync(dev_labels, c->labels_prev(), (size_t)c->n * 4, cudaMemcpyDeviceToDevice, s));
        return 0;
    }
    // the fused small-D kernel keeps no labels: one label-only pass against the centroids of that pass
    return launch_small(c, false, dev_labels, c->updated_since_acc ? c->centroids_prev() : c->centroids(), s);
}

int kmc_labels_to_host(kmc_ctx* c, const int32_t* dev_labels, void* host_labels, int label_bytes, int64_t n_labels,
                       void* stream) {
    if (!c || (n_labels > 0 && (!dev_labels || !host_labels))) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (label_bytes != 4 && label_bytes != 8) KMC_FAIL(KMC_E_INVALID, "label_bytes must be 4 or 8, not %d", label_bytes);
    if (n_labels <= 0) return 0;
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cudaPointerAttributes attr;
    bool dst_pinned = false;
    if (cudaPointerGetAttributes(&attr, host_labels) == cudaSuccess) dst_pinned = (attr.type == cudaMemoryTypeHost);
    cudaGetLastError();
    char* dst = static_cast<char*>(host_labels);
    if (dst_pinned && label_bytes == 4) {
        KMC_CUDA(cudaMemcpyAsync(dst, dev_labels, (size_t)n_labels * 4, cudaMemcpyDeviceToHost, s));
        KMC_CUDA(cudaStreamSynchronize(s));
        return 0;
    }
    KMC_TRY(ensure_staging(c, kLabelChunk, label_bytes == 8));
    const int64_t per = kLabelChunk / label_bytes;        // labels per chunk
    const int64_t n_chunks = (n_labels + per - 1) / per;
    auto chunk_len = [&](int64_t i) { return (i + 1) * per <= n_labels ? per : n_labels - i * per; };
    for (int64_t i = 0; i <= n_chunks; ++i) {
        const int b = (int)(i & 1);
        if (i < n_chunks) {
            // chunk i: (widen on the device,) copy to the pinned buffer -- or straight into a pinned destination
            const int64_t r0 = i * per, nr = chunk_len(i);
            const void* src = dev_labels + r0;
            if (label_bytes == 8) {
                KMC_TRY(launch_widen_labels(c, dev_labels + r0, static_cast<long long*>(c->dev_stage[b]), nr, s));
                c->launches++;
                src = c->dev_stage[b];
            }
            void* to = dst_pinned ? (void*)(dst + (size_t)r0 * label_bytes) : c->pinned[b];
            KMC_CUDA(cudaMemcpyAsync(to, src, (size_t)nr * label_bytes, cudaMemcpyDeviceToHost, s));
            KMC_CUDA(cudaEventRecord(c->pinned_free[b], s));
        }
        if (i > 0 && !dst_pinned) {
            // chunk i-1 has landed: the host copies it out while chunk i is in flight
            const int pb = b ^ 1;
            KMC_CUDA(cudaEventSynchronize(c->pinned_free[pb]));
            parallel_memcpy(dst + (size_t)(i - 1) * per * label_bytes, c->pinned[pb], (size_t)chunk_len(i - 1) * label_bytes);
        }
    }
    KMC_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int kmc_gather_centroids(kmc_ctx* c, const int32_t* dev_labels, float* dev_out, void* stream) {
    if (!c || (c->n > 0 && (!dev_labels || !dev_out))) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (!c->ws || !c->have_centroids) KMC_FAIL(KMC_E_STATE, "no centroids yet");
    KMC_CUDA(cudaSetDevice(c->device));
    return launch_gather_centroids(c, dev_labels, dev_out, static_cast<cudaStream_t>(stream));
}

int kmc_set_param(kmc_ctx* c, int param, double value) {
    if (!c) KMC_FAIL(KMC_E_INVALID, "null context");
    switch (param) {
        case KMC_PARAM_TC_GAMMA:
            if (!(value >= 0.0)) KMC_FAIL(KMC_E_INVALID, "gamma must be >= 0");
            c->tc_gamma = value;
            return 0;
        case KMC_PARAM_TC_TERMS:
            if (value != 0.0 && value != 1.0 && value != 3.0 && value != 16.0)
                KMC_FAIL(KMC_E_INVALID, "tensor terms must be 0, 1, 3 or 16 (1 term, fp16 accumulators)");
            c->tc_terms = (int)value;
            return 0;
        case KMC_PARAM_REFINE_KERNEL:
            if (value != 0.0 && value != 4.0 && value != 8.0) KMC_FAIL(KMC_E_INVALID, "refine kernel must be 0, 4 or 8");
            c->refine_kernel = (int)value;
            return 0;
        case KMC_PARAM_CHUNKS:
            if (value < 1.0 || value > (double)kMaxChunks || value != (double)(int)value)
                KMC_FAIL(KMC_E_INVALID, "row chunks must be an integer in 1..%d", kMaxChunks);
            c->n_chunks = (int)value;
            return 0;
        case KMC_PARAM_TC_TILE_N:
            if (value != 128.0 && value != 256.0) KMC_FAIL(KMC_E_INVALID, "accumulator tiles hold 128 or 256 centroids");
            c->tc_tile_n = (int)value;
            return 0;
        case KMC_PARAM_TC_PAIR:
            c->tc_pair = value != 0.0;
            return 0;
        case KMC_PARAM_TC_ACC16:
            c->tc_acc16 = value != 0.0;
            return 0;
        case KMC_PARAM_INCREMENTAL:
            c->delta_enabled = value != 0.0;
            c->delta_valid = false;
            return 0;
        case KMC_PARAM_DELTA_ATOMIC_ROWS:
            if (value != (double)(int64_t)value) KMC_FAIL(KMC_E_INVALID, "row count must be an integer");
            c->delta_atomic_rows = (int64_t)value;
            return 0;
        case KMC_PARAM_PROFILE:
            c->profile = value != 0.0;
            c->prof_acc_n = c->prof_upd_n = 0;
            return 0;
    }
    KMC_FAIL(KMC_E_INVALID, "unknown parameter %d", param);
}

int kmc_get_phase_ms(kmc_ctx* c, float* out_ms, int n) {
    if (!c || !out_ms || n < 1) KMC_FAIL(KMC_E_INVALID, "bad argument");
    KMC_CUDA(cudaSetDevice(c->device));
    double sum[4] = {0, 0, 0, 0};
    const int na = c->prof_acc_n < kmc_ctx::kProfRing ? c->prof_acc_n : kmc_ctx::kProfRing;
    const int nu = c->prof_upd_n < kmc_ctx::kProfRing ? c->prof_upd_n : kmc_ctx::kProfRing;
    for (int i = 0; i < na; ++i) {
        const int slot = (c->prof_acc_n - 1 - i) % kmc_ctx::kProfRing;
        cudaEvent_t* e = c->prof_ev[slot];
        KMC_CUDA(cudaEventSynchronize(e[3]));
        float ms = 0.f;
        KMC_CUDA(cudaEventElapsedTime(&ms, e[0], e[1])); sum[0] += ms;
        KMC_CUDA(cudaEventElapsedTime(&ms, e[1], e[2])); sum[1] += ms;
        KMC_CUDA(cudaEventElapsedTime(&ms, e[2], e[3])); sum[2] += ms;
    }
    for (int i = 0; i < nu; ++i) {
        const int slot = (c->prof_upd_n - 1 - i) % kmc_ctx::kProfRing;
        cudaEvent_t* e = c->prof_ev[slot];
        KMC_CUDA(cudaEventSynchronize(e[5]));
        float ms = 0.f;
        KMC_CUDA(cudaEventElapsedTime(&ms, e[4], e[5])); sum[3] += ms;
    }
    const float vals[KMC_PHASE_COUNT] = {na ? (float)(sum[0] / na) : 0.f, na ? (float)(sum[1] / na) : 0.f,
                                         na ? (float)(sum[2] / na) : 0.f, nu ? (float)(sum[3] / nu) : 0.f,
                                         (float)na};
    for (int i = 0; i < n && i < KMC_PHASE_COUNT; ++i) out_ms[i] = vals[i];
    c->prof_acc_n = c->prof_upd_n = 0;
    return 0;
}

int kmc_debug_tc_scores(kmc_ctx* c, float* dev_scores, void* stream) {
    KMC_TRY(check_ready(c, true));
    if (c->path != KMC_PATH_TC) KMC_FAIL(KMC_E_UNSUPPORTED, "context is not on the tensor path");
    if (!dev_scores) KMC_FAIL(KMC_E_INVALID, "null scores");
    KMC_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KMC_TRY(ensure_planes(c, s));
    return launch_tc_assign(c, c->labels(), nullptr, &c->stats()->n_amb[0], dev_scores, 0, c->n, c->tc_terms == 16 ? 16 : 3, nullptr, 0u, s);
}

}  // extern "C"
