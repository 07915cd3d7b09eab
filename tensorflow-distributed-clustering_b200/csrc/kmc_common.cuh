// Shared declarations of libkmc_b200 (sm_100a only).  See include/kmc.h for the C ABI.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/kmc.h"

namespace kmc {

void set_error(const char* fmt, ...);

#define KMC_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            kmc::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,          \
                           cudaGetErrorString(e__));                                     \
            return (int)e__;                                                             \
        }                                                                                \
    } while (0)

#define KMC_TRY(call)                                                                    \
    do {                                                                                 \
        int r__ = (call);                                                                \
        if (r__ != 0) return r__;                                                        \
    } while (0)

#define KMC_FAIL(code, ...)                                                              \
    do {                                                                                 \
        kmc::set_error(__VA_ARGS__);                                                     \
        return (code);                                                                   \
    } while (0)

constexpr int kTcBlockM = 128;   // rows per tensor tile (TMEM lanes)
constexpr int kTcBlockN = 256;   // centroids per tensor tile (TMEM columns of one accumulator)
constexpr int kTcChunkK = 64;    // fp16 elements per 128-byte swizzle row
constexpr int kStatSlots = 16;
// score bias S*|c_k|^2 = sum_i kBiasConst[i] * piece_i with fp16 pieces (|piece| <= 65504): range 8.6e9,
// resolution ~2^-10 absolute
constexpr int kBiasPieces = 6;
constexpr int kBiasCols = 16;     // fp16 columns of a bias-plane row (one K=16 MMA step, 32-byte swizzled rows)
__device__ __constant__ const float kBiasConst[kBiasPieces] = {32768.f, 32768.f, 32768.f, 32768.f, 16.f, 0.0078125f};
constexpr float kBiasMax = 2147483648.f;   // the plane is used only while max_k S*|c_k|^2 < 2^31
constexpr int kUpdateMaxBlocks = 1024;   // CTAs of update_kernel (one shift2 slot each)
constexpr int kMaxChunks = 4;    // row chunks of one accumulate pass (assign of chunk c+1 overlaps accumulate of chunk c)

// Device-resident scalars written by kernels (all doubles / 64-bit so one memset clears them).
struct DeviceStats {
    double inertia;        // last accumulate
    double shift2;         // last update
    unsigned long long n_amb[kMaxChunks];   // rows in the refine worklist of each row chunk (last tensor pass)
    unsigned long long n_over[kMaxChunks];  // ... of them handed on to the tile kernel (walk cap reached); cleared with n_amb
    unsigned int x_absmax_bits;       // max |x| over the shard (fp32 bits), set at bind time
    unsigned int c_absmax_bits;       // max |c| after the last update / set_centroids
    unsigned int c_normmax_bits;      // max_k |C'_k| in scaled units (prep kernel)
    unsigned int watchdog;            // non-zero: a tensor-kernel barrier wait timed out
    float sigma_x;                    // power-of-two scale of the point planes
    float sigma_c;                    // power-of-two scale of the centroid planes (x -2 folded)
    unsigned int c_lonormmax_bits;    // max_k |C' - C'_hi| (scaled units): bound for the 1-term kernel
    unsigned int cn_max_bits;         // max_k S*|c_k|^2 (fp32 bits): the bias plane is valid below 2^31
    float rho16;                      // fp16-accumulator tier: its centroid scale / sigma_c (a power of two <= 1)
    unsigned int tc3_flag[kMaxChunks];      // set by the gate: the 3-term kernel redoes this chunk in this pass
    unsigned int tc_skip[kMaxChunks];       // passes for which the chunk goes straight to the 3-term kernel
    unsigned int tc_period[kMaxChunks];     // length of the next skip period (1, 2, 4 ... 16)
    unsigned int tc_probe[kMaxChunks];      // the next pass keeps its 1-term result whatever the statistics say
    // fp16-accumulator tier of the adaptive policy: 1 = off for this chunk (its wider certification margin made the
    // gate fall back to 3 terms; the fp32-accumulator 1-term kernel, which certifies more rows, is the first tier
    // until `t1_ok` more passes stayed with it -- `rearm` of them at first, doubling with every failure of the fp16 tier)
    unsigned int acc16_off[kMaxChunks];
    unsigned int t1_ok[kMaxChunks];
    unsigned int rearm[kMaxChunks];
    // exact distances evaluated / rows handled by the last two refine launches (slot = launch parity): the next
    // launch reads the previous slot to pick the kernel that fits the walk lengths (see refine.cu)
    unsigned long long walk[2][4];          // [slot]: distances evaluated, rows refined, rows handed to the tile kernel, -
    // incremental accumulation (accumulate.cu): rows whose label changed in the last pass, the part of them the
    // changed-row list holds (0 when the pass fell back to a full one), and that decision
    unsigned long long n_changed;
    unsigned long long n_changed_eff;
    unsigned int acc_full;
    unsigned int acc_sorted;          // incremental pass: 1 = the changed rows go through the sort + segmented sum, 0 = through atomics
    unsigned int upd_done;            // CTAs of the running update_kernel that have stored their shift2 slot
};

}  // namespace kmc

struct kmc_ctx {
    int device = 0;
    int sm_count = 0;
    int64_t n = 0, D = 0, K = 0;
    int64_t n_cap = 0;                  // rows the workspace was laid out for (kmc_resize keeps n <= n_cap)
    uint32_t flags = 0;
    int path = 0;                       // KMC_PATH_*
    // tensor-path geometry
    int64_t D_pad = 0, K_pad = 0;
    double tc_gamma = 0.0;
    int tc_terms = 0;                   // KMC_PARAM_TC_TERMS: 0 = adaptive, 1 or 3 = fixed, 16 = 1 term with fp16 accumulators
    bool tc_acc16 = true;               // KMC_PARAM_TC_ACC16: the adaptive policy starts with the fp16-accumulator tier
    int n_chunks = 1;                   // row chunks per accumulate pass (tensor path, large shards)
    cudaStream_t side_stream = nullptr; // refine + accumulation of chunk c run here while chunk c+1 is assigned
    cudaEvent_t chunk_done[kmc::kMaxChunks] = {};
    cudaEvent_t side_done = nullptr, side_go = nullptr;
    cudaEvent_t nbr_ready = nullptr;    // the neighbour lists of the current centroids are complete (side stream)
    bool nbr_pending = false;           // ... and nobody has waited for them on the caller's stream yet
    int refine_kernel = 0;              // KMC_PARAM_REFINE_KERNEL: 0 = picked on the device, 4 or 8 = fixed
    unsigned int refine_launches = 0;   // parity selects the walk-statistics slot of DeviceStats
    // workspace
    char* ws = nullptr;
    int64_t ws_bytes = 0;
    int64_t need_bytes = 0;
    int64_t off_cprev = 0, off_shift_slots = 0, off_small_slots = 0;
    bool updated_since_acc = false;     // kmc_update ran after the last accumulate (centroids_prev() is current)
    int64_t off_centroids = 0, off_partials = 0, off_stats = 0, off_labels = 0, off_labels2 = 0, off_perm = 0,
            off_changed = 0, off_running = 0, off_overflow = 0,
            off_sort = 0, sort_stride = 0, sort_cursor = 0, sort_offsets = 0, sort_segoff = 0,
            off_xhi = 0, off_xlo = 0, off_xnorm = 0, off_cplanes = 0, off_cnorm = 0, off_cbias = 0, off_ambmask = 0, off_cplane16 = 0, off_cbias16 = 0, off_xext = 0,
            off_ccdist = -1, off_nbr = -1;
    const float* points = nullptr;
    // incremental accumulation: two label buffers (this pass / previous pass), running fp64 sums and counts
    bool delta_enabled = true;          // KMC_PARAM_INCREMENTAL
    bool delta_valid = false;           // the running sums match labels_prev()
    bool last_pass_host_full = true;    // the last pass was a full one by the host's decision (no running state yet)
    int lab_cur = 0;                    // which label buffer the next pass writes
    int64_t chg_cap = 0;                // rows the changed-row list holds; more changed rows -> full pass
    int64_t delta_atomic_rows = -1;     // KMC_PARAM_DELTA_ATOMIC_ROWS: up to this many changed rows are applied with atomics (-1: default)
    bool have_centroids = false;
    bool planes_ready = false;
    bool streaming = false;             // kmc_bind_points_streamed: rows arrive in pieces (kmc_accumulate_rows)
    int64_t stream_rows = 0;            // rows delivered and accumulated so far
    // tensor maps (opaque 128-byte blobs, 64-byte aligned)
    alignas(64) CUtensorMap tmap_xhi;
    alignas(64) CUtensorMap tmap_xlo;
    alignas(64) CUtensorMap tmap_c;
    alignas(64) CUtensorMap tmap_bias;
    alignas(64) CUtensorMap tmap_c_half;      // 128-row boxes: the halves of a centroid stage a CTA pair loads
    alignas(64) CUtensorMap tmap_bias_half;
    alignas(64) CUtensorMap tmap_c16, tmap_bias16, tmap_c16_half, tmap_bias16_half, tmap_xext;   // fp16-accumulator tier
    alignas(64) CUtensorMap tmap_c_quarter;   // 64-row boxes: a CTA's half of a 128-centroid stage
    alignas(64) CUtensorMap tmap_bias_quarter;
    int tc_tile_n = 256;                      // KMC_PARAM_TC_TILE_N: centroids per accumulator tile of the pair kernel
    bool tc_pair = true;                      // KMC_PARAM_TC_PAIR: 1-term tensor kernel on CTA pairs (cta_group::2)
    bool tmaps_ready = false;
    // pinned staging
    void* pinned[2] = {nullptr, nullptr};
    cudaEvent_t pinned_free[2] = {nullptr, nullptr};
    void* dev_stage[2] = {nullptr, nullptr};
    int64_t stage_bytes = 0;
    // optional phase timing: ring of event sextuples (acc: e0..e3, update: e4,e5)
    bool profile = false;
    static constexpr int kProfRing = 32;
    cudaEvent_t prof_ev[kProfRing][6] = {};
    int prof_acc_n = 0, prof_upd_n = 0;
    // bookkeeping
    int64_t launches = 0;
    double host_stats[KMC_STAT_COUNT] = {0};

    template <typename T> T* at(int64_t off) const { return reinterpret_cast<T*>(ws + off); }
    float* centroids() const { return at<float>(off_centroids); }
    float* centroids_prev() const { return at<float>(off_cprev); }
    double* partials() const { return at<double>(off_partials); }
    kmc::DeviceStats* stats() const { return at<kmc::DeviceStats>(off_stats); }
    int32_t* labels() const { return at<int32_t>(lab_cur ? off_labels2 : off_labels); }        // written by the next pass
    int32_t* labels_prev() const { return at<int32_t>(lab_cur ? off_labels : off_labels2); }   // labels of the last pass
    void swap_labels() { lab_cur ^= 1; }
    int32_t* perm() const { return at<int32_t>(off_perm); }
    int32_t* changed() const { return at<int32_t>(off_changed); }
    double* running() const { return at<double>(off_running); }      // [K*D sums | K counts | unused | sum |x|^2]
    // per-chunk scratch of the counting sort: hist | cursor | offsets | segoff
    unsigned int* sort_hist(int chunk) const { return at<unsigned int>(off_sort + chunk * sort_stride); }
    unsigned int* sort_cur(int chunk) const { return at<unsigned int>(off_sort + chunk * sort_stride + sort_cursor); }
    unsigned int* sort_off(int chunk) const { return at<unsigned int>(off_sort + chunk * sort_stride + sort_offsets); }
    unsigned int* sort_seg(int chunk) const { return at<unsigned int>(off_sort + chunk * sort_stride + sort_segoff); }
};

namespace kmc {

// ---- kernels (one .cu each) -------------------------------------------------------------------
// stage.cu
int launch_convert_f64(kmc_ctx* c, const double* src, float* dst, int64_t count, cudaStream_t s);
int launch_gather_centroids(kmc_ctx* c, const int32_t* labels, float* out, cudaStream_t s);
int launch_widen_labels(kmc_ctx* c, const int32_t* src, long long* dst, int64_t count, cudaStream_t s);
int launch_absmax(kmc_ctx* c, const float* x, int64_t count, unsigned int* out_bits, cudaStream_t s);
int launch_split_points(kmc_ctx* c, cudaStream_t s);
int launch_split_points_range(kmc_ctx* c, int64_t r0, int64_t r1, bool keep_scale, cudaStream_t s);
// assign_small.cu
bool small_path_supported(int64_t K, int64_t D);
int launch_small(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s);
int launch_small_into(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                      double* partials, bool fused);
int64_t small_max_ctas(int sm_count);
// assign_simt.cu
// `over_stat` (may be null): the number of worklist rows is also added there (walk statistics of the refine)
int launch_simt_assign(kmc_ctx* c, const int32_t* worklist, const unsigned long long* n_work_dev,
                       int64_t n_work_host, int32_t* labels_out, cudaStream_t s,
                       unsigned long long* over_stat = nullptr);
// assign_tc.cu
bool tc_path_supported(int64_t K, int64_t D);
int tc_setup_tensor_maps(kmc_ctx* c);
// rows [r0, r1) (r0 a multiple of 128); uncertified rows are appended to `worklist`, counted in `amb_counter`
// `terms` = 1 or 3 MMA terms; `gate_flag` (may be null): the launch is a no-op unless *gate_flag == gate_want
// `gate2` (may be null): a second device flag of the same kind
int launch_tc_assign(kmc_ctx* c, int32_t* labels_out, int32_t* worklist, unsigned long long* amb_counter,
                     float* scores_out, int64_t r0, int64_t r1, int terms, const unsigned int* gate_flag,
                     unsigned int gate_want, cudaStream_t s, const unsigned int* gate2 = nullptr, unsigned int gate2_want = 0u);
// rows [r0, r1) (r0 a multiple of 32): masks of uncertified rows -> worklist, their number added to *amb_counter
int launch_compact_worklist(kmc_ctx* c, int32_t* worklist, unsigned long long* amb_counter, int64_t r0, int64_t r1,
                            const unsigned int* gate_flag, unsigned int gate_want, cudaStream_t s);
int launch_tc_gate(kmc_ctx* c, int chunk, int64_t rows, bool acc16_policy, cudaStream_t s);
// refine.cu
int launch_ccdist(kmc_ctx* c, cudaStream_t s);
int launch_refine(kmc_ctx* c, int32_t* labels, const int32_t* worklist, const unsigned long long* n_work_dev,
                  int32_t* overflow, unsigned long long* n_over, cudaStream_t s);
// accumulate.cu
// sums / counts / inertia / sum |x|^2 of rows [r0, r1), ADDED to the partial buffer; `chunk` selects the sort
// scratch; `gate` (may be null): device flag, the kernels run only if *gate == 1
int launch_sorted_accumulate(kmc_ctx* c, const int32_t* labels, int64_t r0, int64_t r1, int chunk,
                             const unsigned int* gate, cudaStream_t s);
// accumulation stage of kmc_accumulate (labels of this pass in c->labels()): incremental where possible
int launch_accumulate_stage(kmc_ctx* c, cudaStream_t s);
// update.cu
int launch_update(kmc_ctx* c, cudaStream_t s);
int launch_prep_centroids(kmc_ctx* c, cudaStream_t s);
int launch_centroid_absmax(kmc_ctx* c, cudaStream_t s);

__host__ __device__ inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// Labels of UNCERTIFIED rows between the tensor kernel and the refine may carry a second anchor candidate: the
// fp16-accumulator tier cannot tell which of two columns of one 32-column block holds the best score when their
// fp16 scores are equal, so it stores the first column as the label (bits 0-25) and the second one's position in
// the block, plus one, in bits 26-31.  The refine starts from the nearer of the two and rewrites the label.
constexpr uint32_t kAnchorMask = 0x03FFFFFFu;
__host__ __device__ inline int anchor_first(int32_t lab) { return (int)((uint32_t)lab & kAnchorMask); }
__host__ __device__ inline int anchor_second(int32_t lab) {
    const uint32_t c = (uint32_t)lab >> 26;
    return c ? (int)((((uint32_t)lab & kAnchorMask) & ~31u) + (c - 1u)) : -1;
}

}  // namespace kmc
