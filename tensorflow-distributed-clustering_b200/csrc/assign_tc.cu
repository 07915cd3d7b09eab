// Tensor path of the fused distance + argmin (large K*D): the x.c^T contraction runs on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), operands staged by TMA into
// 128-byte-swizzled shared memory, and the N x K score matrix never leaves the SM.
//
//   score'[i,k] = S*|c_k|^2 + sum_j X'[i,j] * C'[k,j],  X' = sigma_x * x,  C' = -2 sigma_c * c,
//   S = sigma_x*sigma_c  (powers of two), so score' = S * (|c_k|^2 - 2 x_i.c_k) = S*(d - |x|^2).
//
// fp32 accuracy from fp16 tensor inputs: every operand is split at staging time into two fp16
// planes at one scale (v = hi + lo, |lo| <= 2^-11 |hi|).  Three MMAs (hi*hi + lo*hi + hi*lo, NPASS = 3)
// accumulate the fp32-accurate score BASELINE.json asks for (kind::f16 instead of kind::tf32: same ~2^-22
// operand accuracy, half the tensor time per term, and the two planes are exactly the 4 bytes/element of
// the fp32 point).  The default (NPASS = 1) issues hi*hi only and bounds the dropped terms per row; for K > 256 its
// first tier accumulates in fp16 (ACC16: packed TMEM reads and a half2 epilogue, half the ALU work per score; the
// bias MMA step then also adds S|x_i|^2, so the accumulator ends at S|x_i - c_k|^2 where fp16 spacing is fine).
//
// The epilogue finds, per row, the exact best and second-best score of the accumulator values (partition
// epilogue below: 1.44 min/max/logic instructions per score).  Rows whose two best scores are closer than a
// rigorous bound on the score error are flagged in a bit mask (one word per warp and row tile, no atomics
// in this kernel); compact_worklist_kernel turns the masks into a worklist and the exact fp32
// difference-form kernels (refine.cu; assign_simt.cu for K > 8192) re-decide those rows, so the final
// labels equal an exact fp32 argmin with lowest-index ties whatever NPASS was.
//
// Warp roles (384 threads, one CTA per SM, persistent over 128-row tiles):
//   warp 0  TMA producer      warp 1  MMA issuer (one thread)     warp 2  TMEM alloc/dealloc
//   warps 4-11  epilogue: tcgen05.ld 32 lanes x 32 columns (two warps per TMEM lane quarter, each scanning
//   one 128-column half of the accumulator tile; they meet once per row tile on one-way named barriers)
// (cta_group::2: two CTAs on the SMs of one TPC share every MMA, M = 256, each staging half of every centroid stage.)
// Pipelines: A tiles (points, resident for the whole K sweep, double buffered when they fit), B stages
// (256 centroids x 64 columns, 3-4 deep ring; with AUG one 8 KB bias-plane stage per tile adds S*|c_k|^2
// through an extra K=16 MMA), two TMEM accumulators of 256 columns.
//
// Replaces reference scripts/distribuitedClustering.py:221-234 for shapes where K*D makes the
// distance computation a dense contraction.
#include "kmc_common.cuh"
#include "kmc_device.cuh"
#include <type_traits>

namespace kmc {

constexpr uint32_t kTileBytesA = kTcBlockM * 128;   // 128 rows x 128 B  = 16 KB per (plane, chunk)
constexpr uint32_t kTmemCols = 512;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;
constexpr float kPackRel = 1.6e-5f;                 // > 2^-16: relative weight of the 7 packed index bits

// Optional cycle trace of the barrier hand-offs (build with -DKMC_TRACE; tools/trace_run.py reads it): CTA 0
// stamps clock64 -- and %globaltimer at slot 1, which gives the SM clock the kernel really runs at -- for tile
// steps 400..655.  Slots: 0/1 warp 4 before its t_full wait (cycles / ns), 2 t_full seen, 3 t_empty arrive,
// 4 / 5 start / end of the end-of-row-tile section, 6 MMA issuer got the TMEM buffer, 7 issuer committed t_full.
#ifdef KMC_TRACE
__device__ long long g_trace[256][8];
__device__ __forceinline__ long long trace_ns() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define KMC_TRACE_AT(step, k) do { if (blockIdx.x == 0 && (step) >= 400u && (step) < 656u) g_trace[(step) - 400u][k] = clock64(); } while (0)
#define KMC_TRACE_NS(step, k) do { if (blockIdx.x == 0 && (step) >= 400u && (step) < 656u) g_trace[(step) - 400u][k] = trace_ns(); } while (0)
#else
#define KMC_TRACE_AT(step, k) do {} while (0)
#define KMC_TRACE_NS(step, k) do {} while (0)
#endif

struct TcParams {
    int64_t n;
    int K, K_pad;
    int n_mtiles, n_ntiles;   // row tiles of this launch (starting at tile0), centroid tiles
    int tile0;
    unsigned long long* amb_counter;
    const unsigned int* gate_flag;   // null, or: run only if *gate_flag == gate_want (device-side precision policy)
    unsigned int gate_want;
    const unsigned int* gate2;       // null, or a second condition of the same kind
    unsigned int gate2_want;
    const float* xnorm;
    const float* cnorm;
    DeviceStats* stats;
    int32_t* labels;
    int32_t* worklist;   // non-null: uncertified rows are recorded (as bits of amb_mask; compacted afterwards)
    uint32_t* amb_mask;  // one word per 32 rows (word index = row / 32): bit i = row 32*w + i is uncertified
    float* scores;       // debug: [n, K] or nullptr
    float gamma;         // accumulate + split error of the 3-term score, relative to |X'||C'|
    float gamma1;        // accumulate error of the 1-term score
    float gamma16;       // fp16 accumulators: (MMA instructions per tile) * 2^-10
    int use_bias_mma;    // this launch is one of the AUG / plain pair: run only if AUG matches the device-side check
    uint32_t pack_mask;  // 0xFFFFFF80, see pack_idx
    uint32_t blk_mask;   // ~(128 / kEpiG - 1): the packed block index of the partition epilogue
    const float* x;      // fp32 points and centroids, D: exact distances of the rare rows whose best block
    const float* cent;   // holds several cells within the packing truncation (see the partition epilogue)
    int D;
};

// Shared-memory plan.  With one MMA term only the hi plane of the points is staged, which leaves
// room on the SM for the CTAs of the accumulation kernels that run concurrently on the side stream.
// NCTA = 2: a CTA pair shares every MMA (cta_group::2, M = 256): each CTA stages its own 128 rows of points and
// HALF of every centroid stage (128 centroids x 64 columns, 16 KB), so an SM pulls half the centroid bytes
// from L2 and its tensor core reads half the B bytes from shared memory per score.
// BN = centroids per accumulator tile: 256 (two TMEM buffers, the default) or, for CTA pairs, 128 (FOUR buffers,
// KMC_PARAM_TC_TILE_N).  Measured on one B200: four buffers are slower (config 4: 30.2 ms against 23.8 ms per pass,
// config 5 slice: 95.7 against 79.1): twice as many tile steps, each with its fixed hand-off and bookkeeping cost,
// outweigh the deeper MMA / epilogue overlap -- the 256-wide kernel already runs into the board's power cap.
// (One CTA per MMA cannot use N = 128 at all: A + B would need 8 KB of shared-memory reads per 64-cycle MMA.)
template <int KC, int NPASS, int NCTA = 1, int BN = kTcBlockN, bool ACC16 = false> struct TcCfg {
    static constexpr int kPlanes = (NPASS == 1) ? 1 : 2;
    static constexpr int kTBufs = (int)kTmemCols / BN;                // accumulator tiles in TMEM
    static constexpr uint32_t kStageBTotal = BN * 128;                // one stage: BN centroids x 64 columns
    static constexpr uint32_t kBiasStageTotal = BN * kBiasCols * 2;
    static constexpr uint32_t kStageB = kStageBTotal / NCTA;          // bytes of one B stage in THIS CTA
    static constexpr uint32_t kBiasStage = kBiasStageTotal / NCTA;
#ifndef KMC_TC_STAGES1
#define KMC_TC_STAGES1 3
#endif
    static constexpr int kStages = (NCTA == 2) ? (BN == 128 ? 12 : 6) : ((NPASS == 1) ? KMC_TC_STAGES1 : ((KC == 1) ? 4 : 3));
    static constexpr uint32_t kBBytes = kStages * kStageB;
    // 1 term: a constant 128 x 16 fp16 tile (32-byte swizzled) holding kBiasConst in its first columns; multiplied
    // with a stage of the bias plane it makes the MMA add S*|c_k|^2 to the accumulator (no FADD per score in the
    // epilogue)
    // (fp16-accumulator tier: not a constant but the row tile's own 128 x 16 extension, one per points buffer)
    static constexpr uint32_t kAExtTile = kTcBlockM * 32;                           // 128 rows x 32 B
    static constexpr uint32_t kAExtBytes = ACC16 ? 4 * kAExtTile : (NPASS == 1) ? kAExtTile : 0;
    static constexpr uint32_t kOther = kBBytes + kAExtBytes + ((NPASS == 1) ? 8 * 128 * 4 + 1536 : 2 * kTcBlockN * 4) + 512;
    // the points tile is double buffered whenever two copies fit next to the B ring (227 KB per CTA)
    // ... and four deep for one-chunk tiles (D <= 64), whose short K sweep leaves little time per row tile to hide
    // the load latency
    static constexpr int kABufs = (KC == 1 && 4 * kPlanes * KC * kTileBytesA + kOther <= 232448u) ? 4
                                : (2 * kPlanes * KC * kTileBytesA + kOther <= 232448u) ? 2 : 1;
    static constexpr uint32_t kABytes = kABufs * kPlanes * KC * kTileBytesA;
    static constexpr uint32_t kAExtOff = kABytes + kBBytes;
    static constexpr uint32_t kCnOff = kAExtOff + kAExtBytes;
    // 3 terms: 2 x 256 bias floats staged per tile by the whole epilogue; 1 term: 128 floats per epilogue warp.
    // Either region doubles as the buffer that merges the two column halves of a row (384 floats).
    // (with 1 term the merge has its own 1536 bytes behind the per-warp slices, so it needs no block barrier)
    static constexpr uint32_t kCnBytes = (NPASS == 1) ? 8 * 128 * 4 + 1536 : 2 * kTcBlockN * 4;
    static constexpr uint32_t kBarOff = kCnOff + kCnBytes;            // mbarriers
    static constexpr uint32_t kSmemBytes = kBarOff + 512;
};

// Wait on an mbarrier phase.  The single-thread roles (producer, MMA issuer) share their scheduler
// with epilogue warps, so a waiting thread backs off with nanosleep instead of spinning on the issue
// slots the epilogue needs.  SLEEP_NS = 0 for the epilogue's own waits.
#ifndef KMC_WAIT_HINT_NS
#define KMC_WAIT_HINT_NS 4000u
#endif
template <unsigned SLEEP_NS>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, DeviceStats* st, unsigned code) {
    if (mbar_try_wait(bar, parity)) return;
#ifdef KMC_PAIR_SPIN
    if (false) {
#else
    if (SLEEP_NS) {
#endif
        // single-thread roles: parked by the hardware between polls (suspend-time hint), no clock reads
        unsigned spins = 0;
        while (!mbar_try_wait_hint(bar, parity, KMC_WAIT_HINT_NS)) {
            if (++spins > (1u << 22)) {        // >= 80 ms of polling (legitimate waits last microseconds): fail loudly
                st->watchdog = code;
                __threadfence_system();
                __trap();
            }
        }
        return;
    }
    const long long t0 = clock64();
    unsigned spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 1023u) == 0 && clock64() - t0 > 6000000000LL) {   // ~3 s: fail loudly, do not hang
            st->watchdog = code;
            __threadfence_system();
            __trap();
        }
    }
}

// Make the data dependence of the asynchronous TMEM load explicit for the compiler.
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]),
                   "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]),
                   "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]),
                   "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}

// Running best / second-best of one row.  The values carry the column index (within the warp's
// 128-column half tile) in their 7 lowest mantissa bits, so the argmin needs no compare+select
// per element: the min/max chain alone moves the index along.  The scores are offset per row by
// S*|x|^2 (making them S*d >= 0, small for the best centroid), so the 2^-16 relative truncation
// is far below the kernel's own error bound.
struct Top2 {
    float m1, m2;
};

// `mask` (0xFFFFFF80) arrives as a kernel parameter on purpose: with two immediates the compiler
// emits two LOP3 per element, with the mask in a register it is one (a & mask) | j.
__device__ __forceinline__ float pack_idx(float v, int j, uint32_t mask) {
    return __uint_as_float((__float_as_uint(v) & mask) | (uint32_t)j);
}

// two packed scores into a top-2 chain: 5 min/max ops (one of them 3-input) per 2 elements
__device__ __forceinline__ void top2_pair(Top2& t, float p0, float p1) {
    const float a = fminf(p0, p1), b = fmaxf(p0, p1);
    const float u = fmaxf(t.m1, a);
    t.m2 = fminf(fminf(t.m2, u), b);
    t.m1 = fminf(t.m1, a);
}

// 32 accumulator columns of one row (column j0 + 0..31 of the warp's half tile)
template <bool SCORES, bool USE_OFF, bool USE_CN>
__device__ __forceinline__ void epi_chunk(const uint32_t (&r)[32], const float* __restrict__ cn, float off, int j0,
                                          uint32_t mask, Top2& e, Top2& o, float* __restrict__ score_row, int kbase, int K,
                                          float inv_s) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        float v0 = __uint_as_float(r[4 * q + 0]), v1 = __uint_as_float(r[4 * q + 1]);
        float v2 = __uint_as_float(r[4 * q + 2]), v3 = __uint_as_float(r[4 * q + 3]);
        if (USE_CN) {
            const float4 c4 = *reinterpret_cast<const float4*>(cn + 4 * q);
            v0 += c4.x; v1 += c4.y; v2 += c4.z; v3 += c4.w;
        }
        if (USE_OFF) { v0 += off; v1 += off; v2 += off; v3 += off; }
        top2_pair(e, pack_idx(v0, j0 + 4 * q + 0, mask), pack_idx(v1, j0 + 4 * q + 1, mask));
        top2_pair(o, pack_idx(v2, j0 + 4 * q + 2, mask), pack_idx(v3, j0 + 4 * q + 3, mask));
        if (SCORES && score_row) {
            if (kbase + 4 * q + 0 < K) score_row[kbase + 4 * q + 0] = (v0 - off) * inv_s;
            if (kbase + 4 * q + 1 < K) score_row[kbase + 4 * q + 1] = (v1 - off) * inv_s;
            if (kbase + 4 * q + 2 < K) score_row[kbase + 4 * q + 2] = (v2 - off) * inv_s;
            if (kbase + 4 * q + 3 < K) score_row[kbase + 4 * q + 3] = (v3 - off) * inv_s;
        }
    }
}

// ---- partition epilogue ---------------------------------------------------------------------------
// Exact best / second-best of a row WITHOUT carrying an index through every score.  The columns a
// thread scans are covered by two partitions in which any two columns share at most one cell:
//   A: column mod G  -> G running minima over the whole row, raw scores, min3 chains (half an op per score)
//   B: column div G  -> minimum of each block of G consecutive columns (half an op per score); only the
//      block minima carry an index (block number in the low mantissa bits) through an exact top-2 chain.
// The best score m1 is the smallest A cell (cell g*), its block h* comes from the best block minimum.  Every
// other column lies outside cell g* or outside block h*, so
//   second best = min( min_{g != g*} A_g , second-smallest block minimum )            (exact)
// and the column of the best is h*·G + g*.  Cost: (1 + 2.5/G + 1/G) min/max/logic ops per score instead of
// 3.5 (2.875 instead of 7 dispatch cycles at G = 8).  Equal or truncation-close candidates give
// m2 - m1 <= pack error, i.e. an uncertified row that the exact refine decides.
#ifndef KMC_EPI_G
#define KMC_EPI_G 8
#endif
#ifdef KMC_EPI_OLD
constexpr bool kOldEpi = true;
#else
constexpr bool kOldEpi = false;
#endif
constexpr int kEpiG = KMC_EPI_G;                    // 8 or 16
constexpr int kEpiBlocksPerHalf = 128 / kEpiG;      // block numbers packed into the block minima
constexpr float kBlkPackRel = (float)kEpiBlocksPerHalf * 1.3e-7f;   // > 2^-23 * blocks

__device__ __forceinline__ float fmin3(float a, float b, float c) { return fminf(fminf(a, b), c); }

// exact fp32 squared distance (difference form) of one row to one centroid; rare path, one thread per row
__device__ __noinline__ float exact_dist_row(const float* __restrict__ x, const float* __restrict__ c, int D) {
    float acc = 0.f;
    for (int j = 0; j < D; ++j) {
        const float d = x[j] - c[j];
        acc = fmaf(d, d, acc);
    }
    return acc;
}

template <bool USE_OFF, bool USE_CN>
__device__ __forceinline__ void epi_chunk_part(const uint32_t (&r)[32], const float* __restrict__ cn, float off,
                                               int blk0, uint32_t mask, float (&ga)[kEpiG], Top2& tb) {
    float v[32];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        v[4 * q + 0] = __uint_as_float(r[4 * q + 0]); v[4 * q + 1] = __uint_as_float(r[4 * q + 1]);
        v[4 * q + 2] = __uint_as_float(r[4 * q + 2]); v[4 * q + 3] = __uint_as_float(r[4 * q + 3]);
        if (USE_CN) {
            const float4 c4 = *reinterpret_cast<const float4*>(cn + 4 * q);
            v[4 * q + 0] += c4.x; v[4 * q + 1] += c4.y; v[4 * q + 2] += c4.z; v[4 * q + 3] += c4.w;
        }
        if (USE_OFF) { v[4 * q + 0] += off; v[4 * q + 1] += off; v[4 * q + 2] += off; v[4 * q + 3] += off; }
    }
    // partition A
#pragma unroll
    for (int g = 0; g < kEpiG; ++g) {
        if (kEpiG == 8) {
            ga[g] = fmin3(ga[g], v[g], v[8 + g]);
            ga[g] = fmin3(ga[g], v[16 + g], v[24 + g]);
        } else {
            ga[g] = fmin3(ga[g], v[g], v[16 + g]);
        }
    }
    // partition B
    float bm[32 / kEpiG];
#pragma unroll
    for (int b = 0; b < 32 / kEpiG; ++b) {
        const float* w = v + b * kEpiG;
        float m;
        if (kEpiG == 8) {
            m = fminf(fmin3(w[0], w[1], w[2]), fmin3(fmin3(w[3], w[4], w[5]), w[6], w[7]));
        } else {
            const float t0 = fmin3(w[0], w[1], w[2]), t1 = fmin3(w[3], w[4], w[5]), t2 = fmin3(w[6], w[7], w[8]);
            const float t3 = fmin3(w[9], w[10], w[11]), t4 = fmin3(w[12], w[13], w[14]);
            m = fminf(fmin3(t0, t1, t2), fmin3(t3, t4, w[15]));
        }
        bm[b] = pack_idx(m, blk0 + b, mask);
    }
#pragma unroll
    for (int b = 0; b < 32 / kEpiG; b += 2) top2_pair(tb, bm[b], bm[b + 1]);
}

// ---- fp16-accumulator tier (ACC16) ---------------------------------------------------------------
// The same exact partition scheme on packed half2 registers: with fp16 accumulators a tcgen05.ld delivers two
// scores per register and HMNMX2 (one dispatch cycle) handles both, so a score costs ~1.2 dispatch cycles
// instead of 2.9.  Per thread and 64-column load (32 registers):
//   A: register index mod 16, and the half -> 32 running minima (16 half2), one HMNMX2 per register;
//   B: 16 consecutive registers = a block of 32 columns -> a 15-deep... 15-op HMNMX2 tree, the two halves merged
//      in fp32 (the conversion is exact and leaves 13 zero mantissa bits: the 2-bit block number packs without
//      any truncation), then the exact top-2 chain over block minima.
// A cell and a block share exactly one column, so best / second best / column follow as in the fp32 scheme.
constexpr int kH2Cells = 16;                        // half2 registers of partition A (32 cells)
__device__ __forceinline__ __half2 as_h2(uint32_t v) { return *reinterpret_cast<const __half2*>(&v); }
// min of two half2 as ONE HMNMX2 (full rate).  Through __hmin2 the compiler fuses pairs of them into the 3-input
// VHMNMX, which the epilogue measured slower than the two instructions it replaces; the asm statement keeps them apart.
__device__ __forceinline__ uint32_t h2min(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

__device__ __forceinline__ void epi_chunk_h2(const uint32_t (&r)[32], int blk0, uint32_t mask, uint32_t (&ga)[kH2Cells],
                                             Top2& tb) {
#pragma unroll
    for (int j = 0; j < kH2Cells; ++j) {
        ga[j] = h2min(ga[j], r[j]);
        ga[j] = h2min(ga[j], r[kH2Cells + j]);
    }
    float bm[2];
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        const uint32_t* w = r + b * 16;
        const uint32_t t0 = h2min(h2min(w[0], w[1]), h2min(w[2], w[3]));
        const uint32_t t1 = h2min(h2min(w[4], w[5]), h2min(w[6], w[7]));
        const uint32_t t2 = h2min(h2min(w[8], w[9]), h2min(w[10], w[11]));
        const uint32_t t3 = h2min(h2min(w[12], w[13]), h2min(w[14], w[15]));
        const __half2 m = as_h2(h2min(h2min(t0, t1), h2min(t2, t3)));
        bm[b] = pack_idx(fminf(__low2float(m), __high2float(m)), blk0 + b, mask);
    }
    top2_pair(tb, bm[0], bm[1]);
}

template <int KC, bool SCORES, int NPASS, int EW, bool AUG, int NCTA, int BN = kTcBlockN, bool ACC16 = false>   // EW = epilogue warps; AUG = bias through the MMA
#ifndef KMC_TC_MAXNREG
#define KMC_TC_MAXNREG 128
#endif
__global__ void __maxnreg__(SCORES ? 128 : KMC_TC_MAXNREG)
tc_assign_kernel(const __grid_constant__ CUtensorMap tm_xhi, const __grid_constant__ CUtensorMap tm_xlo,
                 const __grid_constant__ CUtensorMap tm_c, const __grid_constant__ CUtensorMap tm_bias,
                 const __grid_constant__ CUtensorMap tm_xext, const TcParams p) {
    using Cfg = TcCfg<KC, NPASS, NCTA, BN, ACC16>;
    static_assert(BN == kTcBlockN || (BN == 128 && NCTA == 2), "128-centroid tiles exist for CTA pairs only");
    constexpr int NB = Cfg::kTBufs;              // accumulator tiles in TMEM (2 or 4)
    // kind::f16, fp16 x fp16 -> fp32, A and B K-major, N = BN, M = 128 (one CTA) or 256 (the MMA a CTA pair shares)
    // ... accumulator (bits 4-5): 1 = fp32, 0 = fp16 (ACC16)
    constexpr uint32_t kAccFmt = ACC16 ? 0u : 1u;
    constexpr uint32_t kInstrDescT = (kAccFmt << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((NCTA * kTcBlockM) >> 4) << 24);
    static_assert(!ACC16 || (NPASS == 1 && AUG), "fp16 accumulators exist for the 1-term kernel with the bias MMA");
    static_assert(NCTA == 1 || (NPASS == 1 && AUG && !SCORES), "the CTA-pair kernel exists for the 1-term AUG variant");
    if (p.gate_flag && *p.gate_flag != p.gate_want) return;     // uniform for the whole grid
    if (p.gate2 && *p.gate2 != p.gate2_want) return;
    // bias through the MMA (AUG): only with one term (shared memory for the extra tile) and while the fp16
    // pieces of the bias plane can represent every S*|c_k|^2.  That is a device-side fact, so the host launches
    // the AUG and the plain variant back to back and exactly one of them runs (uniform for the whole grid).
    constexpr bool aug = AUG;
    if (!ACC16 && p.use_bias_mma && AUG != (__uint_as_float(p.stats->cn_max_bits) < kBiasMax)) return;
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const uint32_t bar0 = smem_base + Cfg::kBarOff;
    auto a_full = [&](int i) { return bar0 + 8u * i; };
    auto a_empty = [&](int i) { return bar0 + 32u + 8u * i; };
    auto b_full = [&](int i) { return bar0 + 64u + 8u * i; };
    auto b_empty = [&](int i) { return bar0 + 64u + 8u * (Cfg::kStages + i); };
    auto t_full = [&](int i) { return bar0 + 64u + 16u * Cfg::kStages + 8u * i; };
    auto t_empty = [&](int i) { return bar0 + 96u + 16u * Cfg::kStages + 8u * i; };
    static_assert(128 + 16 * Cfg::kStages + 8 <= 512 && Cfg::kABufs <= 4 && NB <= 4, "barrier region");
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::kBarOff + 128 + 16 * Cfg::kStages);

    // CTA pair: rank 0 (the leader) issues every MMA.  The "full" barriers of the operand stages and the "TMEM
    // buffer released" barriers are used in the leader only (both CTAs' TMA bytes and both CTAs' epilogue warps
    // arrive there); the "empty" and "accumulator ready" barriers exist in both CTAs and are signalled by the
    // leader's multicast commits.
    const uint32_t cta_rank = (NCTA == 2) ? cluster_ctarank() : 0u;
    const bool leader = cta_rank == 0u;
    if (threadIdx.x == 0) {
        if (smem_base & 1023u) { p.stats->watchdog = 0xA11u; __trap(); }
        for (int i = 0; i < 4; ++i) { mbar_init(a_full(i), 1); mbar_init(a_empty(i), 1); }
        for (int i = 0; i < Cfg::kStages; ++i) { mbar_init(b_full(i), 1); mbar_init(b_empty(i), 1); }
        for (int i = 0; i < NB; ++i) { mbar_init(t_full(i), 1); mbar_init(t_empty(i), EW * NCTA); }
        fence_barrier_init();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_xhi);
        tma_prefetch_desc(&tm_xlo);
        tma_prefetch_desc(&tm_c);
    }
    if (warp == 2) {
        if (NCTA == 2) { tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols); tmem_relinquish_pair(); }
        else { tmem_alloc(smem_u32(tmem_slot), kTmemCols); tmem_relinquish(); }
    }
    if (AUG && !ACC16) {
        // constant A tile of the bias step (128 rows x 16 fp16), in the 32-byte-swizzled K-major layout TMA
        // would produce: row r, 16-byte chunk c lives at r*32 + ((c ^ ((r >> 2) & 1)) << 4); only chunk 0 is non-zero
        uint4* ext = reinterpret_cast<uint4*>(smem + Cfg::kAExtOff);
        for (int i = threadIdx.x; i < kTcBlockM * 2; i += blockDim.x) ext[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncthreads();
        if (threadIdx.x < kTcBlockM) {
            __half* row = reinterpret_cast<__half*>(smem + Cfg::kAExtOff + threadIdx.x * 32 + (((threadIdx.x >> 2) & 1) << 4));
#pragma unroll
            for (int i = 0; i < kBiasPieces; ++i) row[i] = __float2half_rn(kBiasConst[i]);
        }
        fence_proxy_async();       // generic-proxy writes -> visible to the tensor core's async proxy
    }
    tc_fence_before_sync();
    __syncthreads();
    if (NCTA == 2) cluster_sync_all();          // the peer's barriers are initialised before anything arrives on them
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    // work units: 128-row tiles per CTA (NCTA = 1) or 256-row tile pairs per CTA pair (both CTAs of a pair walk
    // the same number of units; the odd tile at the end is an out-of-range tile for rank 1: zero-filled loads,
    // no rows written)
    const int unit = (int)blockIdx.x / NCTA, n_units_grid = (int)gridDim.x / NCTA;
    const int n_units = (p.n_mtiles + NCTA - 1) / NCTA;
    const int n_my_tiles = (n_units - unit + n_units_grid - 1) / n_units_grid;
    auto tile_of = [&](int it) { return p.tile0 + (unit + it * n_units_grid) * NCTA + (int)cta_rank; };
    // When the centroid stages of a whole K sweep fit in the ring (config 3: 256 centroids, one bias stage + one
    // 64-column stage) they are loaded once and stay resident: no reload and no "empty" hand-off per row tile.
    const int stages_per_tile = p.n_ntiles * ((aug ? 1 : 0) + KC * ((NPASS == 1) ? 1 : 2));
    const bool b_resident = stages_per_tile <= Cfg::kStages;

    if (warp == 0) {
        // ===================== TMA producer (whole warp, convergent; elect.sync inside the helpers) =====================
        {
            int pslot = 0;
            uint32_t ppar = 1u;                  // "empty" parity of the ring's current round (fresh barriers pass)
            auto pnext = [&]() { if (++pslot == Cfg::kStages) { pslot = 0; ppar ^= 1u; } };
            // one tile load: own shared memory, bytes counted on the leader's "full" barrier (pair) or our own
            auto load_tile = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1, uint32_t bar, uint64_t policy) {
                if (NCTA == 2) e_tma_load_2d_hint_pair(dst, tm, c0, c1, mapa_cluster(bar, 0u), policy);
                else e_tma_load_2d_hint(dst, tm, c0, c1, bar, policy);
            };
            auto load_a = [&](int it) {
                const int abuf = it % Cfg::kABufs;
                const int m0 = tile_of(it) * kTcBlockM;
                mbar_wait<200>(a_empty(abuf), (((uint32_t)it / Cfg::kABufs) & 1u) ^ 1u, p.stats, 0x100u + abuf);
                if (leader) e_mbar_arrive_expect_tx(a_full(abuf), (uint32_t)NCTA * (Cfg::kPlanes * KC * kTileBytesA + (ACC16 ? Cfg::kAExtTile : 0u)));
#pragma unroll
                for (int pl = 0; pl < Cfg::kPlanes; ++pl)
#pragma unroll
                    for (int kc = 0; kc < KC; ++kc)
                        load_tile(smem_base + ((abuf * Cfg::kPlanes + pl) * KC + kc) * kTileBytesA,
                                  pl == 0 ? &tm_xhi : &tm_xlo, kc * kTcChunkK, m0, a_full(abuf), kEvictFirst);
                if (ACC16)       // the rows' extension for the bias step: constants + pieces of |X'|^2
                    load_tile(smem_base + Cfg::kAExtOff + abuf * Cfg::kAExtTile, &tm_xext, 0, m0, a_full(abuf), kEvictFirst);
            };
            constexpr int kAhead = (Cfg::kABufs > 2) ? Cfg::kABufs - 1 : 1;      // row tiles requested ahead
            for (int j = 0; j < kAhead && j < n_my_tiles; ++j) load_a(j);
            for (int it = 0; it < n_my_tiles; ++it) {
                for (int nt = 0; nt < p.n_ntiles; ++nt) {
                    if (b_resident && it > 0) {
                        if (nt == 0 && it + kAhead < n_my_tiles) load_a(it + kAhead);
                        continue;
                    }
                    auto load_bias_stage = [&]() {      // stage of the bias plane: 256 centroids x 16 columns (8 KB of the slot)
                        const int slot = pslot;
                        mbar_wait<100>(b_empty(slot), ppar, p.stats, 0x280u + slot);
                        if (leader) e_mbar_arrive_expect_tx(b_full(slot), Cfg::kBiasStageTotal);
                        load_tile(smem_base + Cfg::kABytes + slot * Cfg::kStageB, &tm_bias, 0,
                                  nt * BN + (int)cta_rank * (BN / NCTA), b_full(slot), kEvictLast);
                        pnext();
                    };
                    if (aug && !ACC16) load_bias_stage();      // (fp16 accumulators: the bias step comes LAST, see the issuer)
                    for (int kc = 0; kc < KC; ++kc) {
#pragma unroll
                        for (int pl = 0; pl < ((NPASS == 1) ? 1 : 2); ++pl) {
                            const int slot = pslot;
                            mbar_wait<100>(b_empty(slot), ppar, p.stats, 0x200u + slot);
                            if (leader) e_mbar_arrive_expect_tx(b_full(slot), Cfg::kStageBTotal);
                            load_tile(smem_base + Cfg::kABytes + slot * Cfg::kStageB, &tm_c, kc * kTcChunkK,
                                      pl * p.K_pad + nt * BN + (int)cta_rank * (BN / NCTA), b_full(slot),
                                      kEvictLast);
                            pnext();
                        }
                    }
                    if (aug && ACC16) load_bias_stage();
                    // prefetch the next row tile's points as early as its buffer can be free
                    const int pf_after = (Cfg::kABufs >= 2) ? 0 : p.n_ntiles - 1;
                    if (nt == pf_after && it + kAhead < n_my_tiles) load_a(it + kAhead);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (pair: the leader CTA only) =====================
        auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t acc) {
            if (NCTA == 2) e_umma_f16_pair(d, da, db, kInstrDescT, acc); else e_umma_f16(d, da, db, kInstrDescT, acc);
        };
        auto commit = [&](uint32_t bar) {
            if (NCTA == 2) e_umma_commit_pair(bar, (uint16_t)3); else e_umma_commit(bar);
        };
        if (leader) {          // the whole warp, convergent: elect.sync inside mma() / commit() picks the issuing lane
            uint32_t ac = 0;
            // ring position and "full" parity; resident stages: slot = position in the K sweep, and their only
            // phase stays complete (parity 0 for ever)
            int mslot = 0;
            uint32_t mpar = 0u;
            const int mwrap = b_resident ? stages_per_tile : Cfg::kStages;
            const uint32_t mflip = b_resident ? 0u : 1u;
            auto mnext = [&]() { if (++mslot == mwrap) { mslot = 0; mpar ^= mflip; } };
            const uint64_t b_desc0 = umma_desc_sw128(smem_base + Cfg::kABytes);
            const uint64_t bias_desc0 = umma_desc_sw32(smem_base + Cfg::kABytes);
            const uint64_t ext_desc = umma_desc_sw32(smem_base + Cfg::kAExtOff);
            for (int it = 0; it < n_my_tiles; ++it) {
                const int abuf = it % Cfg::kABufs;
                mbar_wait<40>(a_full(abuf), ((uint32_t)it / Cfg::kABufs) & 1u, p.stats, 0x300u + abuf);
                // operand descriptors differ in their start-address field only (bytes >> 4, the low bits of the
                // 64-bit word): the per-step descriptors below are base + constant, one add each in the issuing thread
                const uint64_t a_hi = umma_desc_sw128(smem_base + (abuf * Cfg::kPlanes + 0) * KC * kTileBytesA);
                const uint64_t a_lo = umma_desc_sw128(smem_base + (abuf * Cfg::kPlanes + (Cfg::kPlanes - 1)) * KC * kTileBytesA);
                for (int nt = 0; nt < p.n_ntiles; ++nt, ++ac) {
                    const uint32_t buf = ac % (uint32_t)NB;
                    mbar_wait<40>(t_empty(buf), ((ac / (uint32_t)NB) & 1u) ^ 1u, p.stats, 0x400u + buf);
                    KMC_TRACE_AT(ac, 6);
                    tc_fence_after_sync();
                    const uint32_t d_tmem = tmem_base + buf * BN;
                    uint32_t accum = 0;
                    // bias step (one k-step of 16): constants x pieces of S|c_k|^2 -- and, with fp16 accumulators,
                    // pieces of |X'_i|^2 x constants, so that the accumulator ends at S16 |x_i - c_k|^2.  With fp16
                    // accumulators it comes LAST: until then the accumulator holds partial dot products only, the
                    // smallest magnitudes (and fp16 rounding steps) the sum can go through.
                    auto bias_step = [&](uint64_t a_ext_desc) {
                        const int slot = mslot;
                        mbar_wait<20>(b_full(slot), mpar, p.stats, 0x580u + slot);
                        tc_fence_after_sync();
                        mma(d_tmem, a_ext_desc, bias_desc0 + (uint64_t)(slot * (int)(Cfg::kStageB >> 4)), accum);
                        accum = 1;
                        if (!b_resident) commit(b_empty(slot));
                        mnext();
                    };
                    if (aug && !ACC16) bias_step(ext_desc);
                    for (int kc = 0; kc < KC; ++kc) {
                        // stage with the centroid hi plane: x_hi*c_hi, then x_lo*c_hi
                        int slot = mslot;
                        mbar_wait<20>(b_full(slot), mpar, p.stats, 0x500u + slot);
                        tc_fence_after_sync();
                        uint64_t b_desc = b_desc0 + (uint64_t)(slot * (int)(Cfg::kStageB >> 4));
                        const uint64_t a_kc = a_hi + (uint64_t)(kc * (int)(kTileBytesA >> 4));
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4) {
                            mma(d_tmem, a_kc + (uint64_t)(2 * k4), b_desc + (uint64_t)(2 * k4), accum);
                            accum = 1;
                        }
                        if (NPASS == 3) {
#pragma unroll
                            for (int k4 = 0; k4 < 4; ++k4)
                                mma(d_tmem, a_lo + (uint64_t)(kc * (int)(kTileBytesA >> 4) + 2 * k4), b_desc + (uint64_t)(2 * k4), 1);
                        }
                        if (!b_resident) commit(b_empty(slot));
                        mnext();
                        if (NPASS == 1) continue;
                        // stage with the centroid lo plane: x_hi*c_lo
                        slot = mslot;
                        mbar_wait<20>(b_full(slot), mpar, p.stats, 0x600u + slot);
                        tc_fence_after_sync();
                        b_desc = b_desc0 + (uint64_t)(slot * (int)(Cfg::kStageB >> 4));
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4)
                            mma(d_tmem, a_kc + (uint64_t)(2 * k4), b_desc + (uint64_t)(2 * k4), 1);
                        if (!b_resident) commit(b_empty(slot));
                        mnext();
                    }
                    if (aug && ACC16) bias_step(umma_desc_sw32(smem_base + Cfg::kAExtOff + abuf * Cfg::kAExtTile));
                    commit(t_full(buf));
                    KMC_TRACE_AT(ac, 7);
                }
                commit(a_empty(abuf));
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: running top-2 per row =====================
        // warps 4-7 scan columns 0..127 of every accumulator tile, warps 8-11 columns 128..255; both
        // groups own the TMEM lane quarter (warp % 4).  Two warps per scheduler hide the TMEM-load
        // and min/max-chain latencies of each other.
        constexpr int kGroups = EW / 4;                 // column groups of the 256-column tile
        constexpr int kCols = BN / kGroups;             // columns scanned per warp and tile (128 or 64)
        constexpr int kEpiThreads = EW * 32;
        const int q = warp & 3;                         // TMEM lane quarter == warp_id % 4
        const int half = (warp - 4) >> 2;               // column group of this warp
        const int et = threadIdx.x - 128;               // 0..kEpiThreads-1
        float* cn_s = reinterpret_cast<float*>(smem + Cfg::kCnOff);
        const float sigma_x = p.stats->sigma_x, sigma_c = p.stats->sigma_c;
        const float inv_s = 1.f / (sigma_x * sigma_c);
        const float cmax = __uint_as_float(p.stats->c_normmax_bits);
        const float clomax = __uint_as_float(p.stats->c_lonormmax_bits);
        // fp16-accumulator tier: its scores are rho16 times those of the fp32 kernels (a smaller power-of-two centroid
        // scale keeps every partial sum inside the fp16 range)
        const float rho16 = ACC16 ? p.stats->rho16 : 1.f;
        const float inv_s16 = inv_s / rho16;
        (void)inv_s16;
        uint32_t ac = 0;
        // row norms of the NEXT row tile are fetched while the current one is scanned
        auto tile_row = [&](int it) { return (int64_t)tile_of(it) * kTcBlockM + q * 32 + lane; };
        // "TMEM buffer released": the leader's barrier, also for the warps of the peer CTA
        const uint32_t t_empty_c0 = (NCTA == 2) ? mapa_cluster(t_empty(0), 0u) : t_empty(0);
        float2 nn_next = make_float2(0.f, 0.f);
        if (n_my_tiles > 0 && tile_row(0) < p.n) nn_next = *reinterpret_cast<const float2*>(p.xnorm + 2 * tile_row(0));
        for (int it = 0; it < n_my_tiles; ++it) {
            const int64_t row = tile_row(it);
            const bool row_ok = row < p.n;
            const float xn = nn_next.x, xlo = nn_next.y;
            nn_next = make_float2(0.f, 0.f);
            if (it + 1 < n_my_tiles && tile_row(it + 1) < p.n)
                nn_next = *reinterpret_cast<const float2*>(p.xnorm + 2 * tile_row(it + 1));
            // certified if the two best scores are further apart than twice the score error bound:
            // 3 terms: operand split + accumulator truncation; 1 term: the dropped x_lo.c and x.c_lo too
            // fp16 accumulators: every MMA instruction of the dot product rounds the accumulator to fp16 (2^-11 of its
            // magnitude, which is at most rho16 * |X'_i| max_k |C'_k| while only partial dot products have been added),
            // on top of the dropped low-order terms of the 1-term kernel
            // (the last MMA step adds the large terms and rounds a small result: its share is the |m1| + |m2| term at
            // the certification below)
            const float thr = ACC16 ? 2.f * rho16 * (xlo * cmax + xn * clomax + p.gamma16 * xn * cmax)
                            : (NPASS == 3) ? 2.f * p.gamma * xn * cmax
                                           : 2.f * (xlo * cmax + xn * clomax + p.gamma1 * xn * cmax);
            // 3 terms: offset by ~S*|x|^2 so that scores are S*d >= 0 and the index packing truncates
            // 2^-16 of a SMALL number; with 1 term the threshold is far above that truncation anyway
            const float off = (NPASS == 3) ? xn * xn * (sigma_c / sigma_x) : 0.f;
            float* score_row = (SCORES && p.scores && row_ok) ? p.scores + row * (int64_t)p.K : nullptr;
            Top2 te = {INFINITY, INFINITY}, to = {INFINITY, INFINITY};     // SCORES (debug) build: per-score index
            int a_e = 0, a_o = 0;
            float ga[kEpiG];                                                // partition epilogue
#pragma unroll
            for (int g = 0; g < kEpiG; ++g) ga[g] = INFINITY;
            uint32_t gh[kH2Cells];                                          // ... of the fp16-accumulator tier (half2 bits)
#pragma unroll
            for (int g = 0; g < kH2Cells; ++g) gh[g] = 0x7C007C00u;         // (+inf, +inf)
            Top2 tb = {INFINITY, INFINITY};
            int a_blk = 0;
            for (int nt = 0; nt < p.n_ntiles; ++nt, ++ac) {
                const uint32_t buf = ac % (uint32_t)NB;
                const float* cn = cn_s;
                if (aug) {
                    // the MMA already added the biases: nothing to stage
                } else if (NPASS == 1) {
                    // one MMA term: every warp stages its own slice of the biases and needs no block-wide
                    // barrier per tile (the warps drift apart freely)
                    float* cw = cn_s + (warp - 4) * kCols;
                    if (lane < kCols / 4)
                        *reinterpret_cast<float4*>(cw + 4 * lane) =
                            *reinterpret_cast<const float4*>(p.cnorm + nt * BN + half * kCols + 4 * lane);
                    __syncwarp();
                    cn = cw;
                } else {
                    float* cw = cn_s + buf * BN;
                    if (et < BN) cw[et] = p.cnorm[nt * BN + et];
                    asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
                    cn = cw + half * kCols;
                }
                if (threadIdx.x == 128) { KMC_TRACE_AT(ac, 0); KMC_TRACE_NS(ac, 1); }
                mbar_wait<0>(t_full(buf), (ac / (uint32_t)NB) & 1u, p.stats, 0x700u + buf);
                if (threadIdx.x == 128) KMC_TRACE_AT(ac, 2);
                tc_fence_after_sync();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN + half * kCols;
                const float e_in = te.m1, o_in = to.m1, b_in = tb.m1;
                const int kbase = nt * BN + half * kCols;
                uint32_t ra[32], rb[32];
                if constexpr (ACC16) {
                    static_assert(!ACC16 || kCols == 128, "two packed loads of 64 columns per warp and tile");
                    tmem_ld_32x64_pack16(taddr, ra);
                    tmem_ld_wait_dep(ra);
                    tmem_ld_32x64_pack16(taddr + 64, rb);
                    if constexpr (SCORES) {
                        if (score_row) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                if (kbase + 2 * i < p.K) score_row[kbase + 2 * i] = __low2float(as_h2(ra[i])) * inv_s16;
                                if (kbase + 2 * i + 1 < p.K) score_row[kbase + 2 * i + 1] = __high2float(as_h2(ra[i])) * inv_s16;
                            }
                        }
                    }
                    epi_chunk_h2(ra, 0, 0xFFFFFFFCu, gh, tb);
                    tmem_ld_wait_dep(rb);
                    if constexpr (SCORES) {
                        if (score_row) {
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                if (kbase + 64 + 2 * i < p.K) score_row[kbase + 64 + 2 * i] = __low2float(as_h2(rb[i])) * inv_s16;
                                if (kbase + 64 + 2 * i + 1 < p.K) score_row[kbase + 64 + 2 * i + 1] = __high2float(as_h2(rb[i])) * inv_s16;
                            }
                        }
                    }
                    epi_chunk_h2(rb, 2, 0xFFFFFFFCu, gh, tb);
                } else {
                tmem_ld_32x32(taddr, ra);
#pragma unroll
                for (int cc = 0; cc < kCols / 32; cc += 2) {
                    tmem_ld_wait_dep(ra);
                    tmem_ld_32x32(taddr + (cc + 1) * 32, rb);
                    if constexpr (SCORES || kOldEpi)
                        epi_chunk<SCORES, NPASS == 3, !AUG>(ra, cn + cc * 32, off, cc * 32, p.pack_mask, te, to, score_row,
                                                            kbase + cc * 32, p.K, inv_s);
                    else
                        epi_chunk_part<NPASS == 3, !AUG>(ra, cn + cc * 32, off, cc * 32 / kEpiG, p.blk_mask, ga, tb);
                    tmem_ld_wait_dep(rb);
                    if (cc + 2 < kCols / 32) tmem_ld_32x32(taddr + (cc + 2) * 32, ra);
                    if constexpr (SCORES || kOldEpi)
                        epi_chunk<SCORES, NPASS == 3, !AUG>(rb, cn + (cc + 1) * 32, off, (cc + 1) * 32, p.pack_mask, te, to,
                                                            score_row, kbase + (cc + 1) * 32, p.K, inv_s);
                    else
                        epi_chunk_part<NPASS == 3, !AUG>(rb, cn + (cc + 1) * 32, off, (cc + 1) * 32 / kEpiG, p.blk_mask,
                                                         ga, tb);
                }
                }
                tc_fence_before_sync();
                __syncwarp();
                if (threadIdx.x == 128) KMC_TRACE_AT(ac, 3);
                if (lane == 0) {
                    if (NCTA == 2 && !leader) mbar_arrive_cluster(t_empty_c0 + 8u * buf); else mbar_arrive(t_empty(buf));
                }
                // a changed best value carries its column in the low bits
                if (__float_as_uint(te.m1) != __float_as_uint(e_in)) a_e = kbase + (int)(__float_as_uint(te.m1) & 0x7Fu);
                if (__float_as_uint(to.m1) != __float_as_uint(o_in)) a_o = kbase + (int)(__float_as_uint(to.m1) & 0x7Fu);
                if (__float_as_uint(tb.m1) != __float_as_uint(b_in))
                    a_blk = ACC16 ? kbase + (int)(__float_as_uint(tb.m1) & 3u) * 32
                                  : kbase + (int)(__float_as_uint(tb.m1) & (uint32_t)(kEpiBlocksPerHalf - 1)) * kEpiG;
            }
            if (threadIdx.x == 128) KMC_TRACE_AT((uint32_t)(it + 1) * (uint32_t)p.n_ntiles - 1u, 4);
            float m1, m2;
            int a1;
            if constexpr (ACC16) {
                // 32 cells (cell g = low half of register g, cell 16+g = its high half), all reductions as half2
                // trees: best = smallest cell; the cells that hold it; second best = the smaller of the best OTHER
                // cell and the second-smallest block minimum (= the best again if two cells tie)
                uint32_t t8[8];
#pragma unroll
                for (int g = 0; g < 8; ++g) t8[g] = h2min(gh[g], gh[8 + g]);
                uint32_t tm = h2min(h2min(h2min(t8[0], t8[1]), h2min(t8[2], t8[3])), h2min(h2min(t8[4], t8[5]), h2min(t8[6], t8[7])));
                const __half m1h = __hmin(__low2half(as_h2(tm)), __high2half(as_h2(tm)));
                m1 = __half2float(m1h);
                const __half2 m1h2 = __half2half2(m1h);
                uint32_t hits = 0, ex[kH2Cells];
#pragma unroll
                for (int g = 0; g < kH2Cells; ++g) {
                    const uint32_t eq = __heq2_mask(as_h2(gh[g]), m1h2);          // 0xFFFF per equal half
                    hits |= (eq & 0x00010001u) << g;
                    ex[g] = gh[g] | (eq & 0x7E007E00u);                           // equal halves -> NaN: ignored by min
                }
#pragma unroll
                for (int g = 0; g < 8; ++g) t8[g] = h2min(ex[g], ex[8 + g]);
                tm = h2min(h2min(h2min(t8[0], t8[1]), h2min(t8[2], t8[3])), h2min(h2min(t8[4], t8[5]), h2min(t8[6], t8[7])));
                const float others = fminf(__low2float(as_h2(tm)), __high2float(as_h2(tm)));   // NaN (no other cell) -> ignored
                const bool tie = (hits & (hits - 1u)) != 0u;
                m2 = tie ? m1 : fminf(tb.m2, others);
                // column of the best: block of the smallest block minimum (its number sits in two mantissa bits that
                // an fp16 value leaves zero) and the cell that holds the best value.  Two such cells (equal fp16
                // scores): which of them lies in that block cannot be told here, so the label carries both columns
                // and the refine, which gets every such row (m2 = m1), starts from the nearer one.
                auto col_of = [](int c) { return c < 16 ? 2 * c : 2 * (c - 16) + 1; };
                const int c1 = hits ? __ffs(hits) - 1 : 0;
                a1 = a_blk + col_of(c1);
                const uint32_t rest = hits & (hits - 1u);
                if (tie) a1 |= (col_of(__ffs(rest) - 1) + 1) << 26;
                // three or more cells with the best value (4 rows in 10^4 at config 4): the label has room for two
                // candidates, so the warp settles these rows here with exact distances of the tied columns
                unsigned need = __ballot_sync(0xffffffffu, (rest & (rest - 1u)) != 0u && row_ok);
                while (need) {
                    const int src = __ffs(need) - 1;
                    need &= need - 1u;
                    const int64_t r_s = __shfl_sync(0xffffffffu, row, src);
                    const uint32_t h_s = __shfl_sync(0xffffffffu, hits, src);
                    const int b_s = __shfl_sync(0xffffffffu, a_blk, src);
                    float bd = INFINITY;
                    int bkk = b_s;
                    for (uint32_t h = h_s; h; h &= h - 1u) {
                        const int k = b_s + col_of(__ffs(h) - 1);
                        if (k >= p.K) continue;
                        float acc = 0.f;
                        for (int j = lane; j < p.D; j += 32) {
                            const float dd = p.x[r_s * p.D + j] - p.cent[(int64_t)k * p.D + j];
                            acc = fmaf(dd, dd, acc);
                        }
                        acc = warp_sum(acc);
                        if (acc < bd) { bd = acc; bkk = k; }
                    }
                    if (lane == src) a1 = bkk;
                }
            } else if constexpr (SCORES || kOldEpi) {
                // merge the two chains of this thread ...
                if (to.m1 < te.m1) { m1 = to.m1; a1 = a_o; m2 = fminf(te.m1, to.m2); }
                else { m1 = te.m1; a1 = a_e; m2 = fminf(to.m1, te.m2); }
            } else {
                // best = smallest A cell (lowest cell on ties); second best = the smaller of the best other A
                // cell and the second-smallest block minimum
                m1 = ga[0];
#pragma unroll
                for (int g = 1; g < kEpiG; ++g) m1 = fminf(m1, ga[g]);
                int gr = 0;
#pragma unroll
                for (int g = kEpiG - 1; g >= 0; --g) gr = (ga[g] == m1) ? g : gr;
                m2 = tb.m2;
#pragma unroll
                for (int g = 0; g < kEpiG; ++g) m2 = fminf(m2, (g == gr) ? INFINITY : ga[g]);
                // Column of the best: block h' of the smallest PACKED block minimum, and the A cell whose
                // minimum equals it up to the packing truncation.  The cell of that block's minimum always
                // matches; when it is the only match the column is exact.  Several matches (an uncertified
                // near tie, ~1e-4 of the rows) are resolved by exact distances of the matching columns of block
                // h', so the refine never starts from a far-away centroid.
                const uint32_t tbits = __float_as_uint(tb.m1) & p.blk_mask;
                uint32_t hits = 0;
#pragma unroll
                for (int g = 0; g < kEpiG; ++g)
                    hits |= ((__float_as_uint(ga[g]) & p.blk_mask) == tbits) ? (1u << g) : 0u;
                a1 = a_blk + (hits ? __ffs(hits) - 1 : gr);
                if ((hits & (hits - 1u)) != 0u && row_ok) {
                    float bd = INFINITY;
                    for (uint32_t h = hits; h; h &= h - 1u) {
                        const int k = a_blk + __ffs(h) - 1;
                        if (k < p.K) {
                            const float d = exact_dist_row(p.x + row * p.D, p.cent + (int64_t)k * p.D, p.D);
                            if (d < bd) { bd = d; a1 = k; }
                        }
                    }
                }
            }
            // ... and the column groups of the row through shared memory
            if (NPASS == 1 && kGroups == 2) {
                // only the two warps that share a TMEM lane quarter meet, on their own named barrier
                // Barrier 2+q: "half 1's result is in mg" (half 1 arrives without waiting, half 0 waits);
                // barrier 6+q: "mg has been read" (half 0 arrives, half 1 waits for it only before its NEXT
                // write).  Neither half ever waits for the other's end-of-tile work.
                float* mg = cn_s + 8 * 128 + q * 96;
                if (half == 1) {
                    if (it > 0) asm volatile("bar.sync %0, 64;" ::"r"(6 + q) : "memory");
                    mg[lane] = m1; mg[32 + lane] = m2; mg[64 + lane] = __int_as_float(a1);
                    asm volatile("bar.arrive %0, 64;" ::"r"(2 + q) : "memory");
                } else {
                    asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                    const float o1 = mg[lane], o2 = mg[32 + lane];
                    const int oa = __float_as_int(mg[64 + lane]);
                    if (it + 1 < n_my_tiles) asm volatile("bar.arrive %0, 64;" ::"r"(6 + q) : "memory");
                    if (o1 < m1) { m2 = fminf(m1, o2); m1 = o1; a1 = oa; }
                    else { m2 = fminf(m2, o1); }
                }
            }
            float* mg = cn_s;                 // 3 terms: the bias buffer is idle here
#pragma unroll
            for (int g = 1; g < ((NPASS == 1 && kGroups == 2) ? 1 : kGroups); ++g) {
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
                if (half == g) {
                    const int r = q * 32 + lane;
                    mg[r] = m1; mg[128 + r] = m2; mg[256 + r] = __int_as_float(a1);
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
                if (half == 0) {
                    const int r = q * 32 + lane;
                    const float o1 = mg[r], o2 = mg[128 + r];
                    const int oa = __float_as_int(mg[256 + r]);
                    if (o1 < m1) { m2 = fminf(m1, o2); m1 = o1; a1 = oa; }
                    else { m2 = fminf(m2, o1); }
                }
            }
            if (half == 0) {                                 // warp-uniform
                bool amb = false;
                if (row_ok) {
                    if ((int)((uint32_t)a1 & kAnchorMask) >= p.K) a1 = 0;
                    if (!p.worklist) a1 = (int)((uint32_t)a1 & kAnchorMask);      // no refine behind this launch: plain labels
                    p.labels[row] = a1;
                    const float pack_err = (ACC16 ? 6.2e-4f : (SCORES || kOldEpi) ? kPackRel : kBlkPackRel) * (fabsf(m1) + fabsf(m2));   // fp16: 1.25 * 2^-11
                    float thr_row = thr;
                    if (NPASS == 1 && !ACC16) {
                        // Tighter bound on the dropped x_lo.(c_k - c_b): |c_k - c_b| <= |x - c_k| + |x - c_b|, and the
                        // distances follow from the scores themselves (P = score + S|x|^2 = S*d, each known to
                        // +-thr/2).  The right-hand side grows more slowly in the score of k than the gap does, so
                        // holding for the second best it holds for every k.  The x.(c_lo,k - c_lo,b) part and the
                        // accumulator rounding stay as they are.
                        const float sx2 = xn * xn * (sigma_c / sigma_x);          // S*|x|^2, rounded up
                        const float rho = 2.f * xlo * sqrtf(sigma_c / sigma_x);
                        const float r1 = sqrtf(fmaxf(m1 + sx2 + 0.5f * thr, 0.f));
                        const float r2 = sqrtf(fmaxf(m2 + sx2 + 0.5f * thr, 0.f));
                        const float tight = rho * (r1 + r2) * 1.001f + 2.f * (xn * clomax + p.gamma1 * xn * cmax);
                        if (r2 > rho) thr_row = fminf(thr, tight);                // (monotonicity needs r > rho / 2)
                    }
                    amb = p.worklist && !((m2 - m1) > thr_row + pack_err);        // NaNs count as uncertified
                }
                // uncertified rows: one 32-bit mask per warp and row tile, no atomic and no dependent store on
                // this path (a returning atomic here costs ~800 cycles per row tile in warps that hold two of the
                // eight TMEM-buffer releases); compact_worklist_kernel turns the masks into the worklist
                const unsigned mask = __ballot_sync(0xffffffffu, amb);
                if (p.worklist && lane == 0 && (row - lane) < p.n) p.amb_mask[(row - lane) >> 5] = mask;
                if (threadIdx.x == 128) KMC_TRACE_AT((uint32_t)(it + 1) * (uint32_t)p.n_ntiles - 1u, 5);
            }
            if (!(NPASS == 1 && kGroups == 2)) asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (NCTA == 2) cluster_sync_all();          // no CTA leaves (or frees TMEM) while its peer may still signal it
    if (warp == 2) {
        tc_fence_after_sync();
        if (NCTA == 2) tmem_dealloc_pair(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
    }
}

// -------------------------------------------------------------------------------------------------
bool tc_path_supported(int64_t K, int64_t D) {
    const int64_t D_pad = round_up(D, kTcChunkK);
    return D >= 16 && D_pad <= 4 * kTcChunkK && K >= 32 && K <= (1 << 20);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int make_map(EncodeTiledFn fn, CUtensorMap* m, void* base, uint64_t cols, uint64_t rows, uint32_t box_rows,
                    uint32_t box_cols = kTcChunkK, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) KMC_FAIL(KMC_E_DEVICE, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

int tc_setup_tensor_maps(kmc_ctx* c) {
    void* fnp = nullptr;
    cudaDriverEntryPointQueryResult qres;
    KMC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &qres));
    if (!fnp || qres != cudaDriverEntryPointSuccess)
        KMC_FAIL(KMC_E_DEVICE, "driver entry point cuTensorMapEncodeTiled not available");
    EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(fnp);
    KMC_TRY(make_map(fn, &c->tmap_xhi, c->at<void>(c->off_xhi), (uint64_t)c->D_pad, (uint64_t)c->n, kTcBlockM));
    KMC_TRY(make_map(fn, &c->tmap_xlo, c->at<void>(c->off_xlo), (uint64_t)c->D_pad, (uint64_t)c->n, kTcBlockM));
    KMC_TRY(make_map(fn, &c->tmap_c, c->at<void>(c->off_cplanes), (uint64_t)c->D_pad, (uint64_t)(2 * c->K_pad),
                     kTcBlockN));
    KMC_TRY(make_map(fn, &c->tmap_bias, c->at<void>(c->off_cbias), (uint64_t)kBiasCols, (uint64_t)c->K_pad, kTcBlockN, kBiasCols,
                     CU_TENSOR_MAP_SWIZZLE_32B));
    // CTA pairs: every CTA loads its half (128 centroids) of a stage
    KMC_TRY(make_map(fn, &c->tmap_c_half, c->at<void>(c->off_cplanes), (uint64_t)c->D_pad, (uint64_t)(2 * c->K_pad),
                     kTcBlockN / 2));
    KMC_TRY(make_map(fn, &c->tmap_bias_half, c->at<void>(c->off_cbias), (uint64_t)kBiasCols, (uint64_t)c->K_pad, kTcBlockN / 2,
                     kBiasCols, CU_TENSOR_MAP_SWIZZLE_32B));
    KMC_TRY(make_map(fn, &c->tmap_xext, c->at<void>(c->off_xext), (uint64_t)16, (uint64_t)c->n, kTcBlockM, 16, CU_TENSOR_MAP_SWIZZLE_32B));
    // fp16-accumulator tier: one centroid plane and a bias plane at their own scale
    KMC_TRY(make_map(fn, &c->tmap_c16, c->at<void>(c->off_cplane16), (uint64_t)c->D_pad, (uint64_t)c->K_pad, kTcBlockN));
    KMC_TRY(make_map(fn, &c->tmap_bias16, c->at<void>(c->off_cbias16), (uint64_t)kBiasCols, (uint64_t)c->K_pad, kTcBlockN, kBiasCols,
                     CU_TENSOR_MAP_SWIZZLE_32B));
    KMC_TRY(make_map(fn, &c->tmap_c16_half, c->at<void>(c->off_cplane16), (uint64_t)c->D_pad, (uint64_t)c->K_pad, kTcBlockN / 2));
    KMC_TRY(make_map(fn, &c->tmap_bias16_half, c->at<void>(c->off_cbias16), (uint64_t)kBiasCols, (uint64_t)c->K_pad, kTcBlockN / 2,
                     kBiasCols, CU_TENSOR_MAP_SWIZZLE_32B));
    // ... or its half of a 128-centroid stage (four accumulator tiles)
    KMC_TRY(make_map(fn, &c->tmap_c_quarter, c->at<void>(c->off_cplanes), (uint64_t)c->D_pad, (uint64_t)(2 * c->K_pad),
                     kTcBlockN / 4));
    KMC_TRY(make_map(fn, &c->tmap_bias_quarter, c->at<void>(c->off_cbias), (uint64_t)kBiasCols, (uint64_t)c->K_pad,
                     kTcBlockN / 4, kBiasCols, CU_TENSOR_MAP_SWIZZLE_32B));
    c->tmaps_ready = true;
    return 0;
}

// Worklist of the rows [32*w0, 32*(w0+nw)) from the masks the tensor kernel wrote: a warp takes 32 mask words,
// reserves the slots of all their set bits with one atomic and writes the row numbers in ascending order
// (consecutive worklist entries are rows close together in memory: the refine kernels gather whole rows).
__global__ void __launch_bounds__(256) compact_worklist_kernel(const uint32_t* __restrict__ mask, int64_t w0, int64_t nw,
                                                               int32_t* __restrict__ worklist,
                                                               unsigned long long* __restrict__ counter,
                                                               const unsigned int* __restrict__ gate_flag,
                                                               unsigned int gate_want) {
    if (gate_flag && *gate_flag != gate_want) return;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t base = warp * 32; base < nw; base += nwarps * 32) {
        const int64_t w = base + lane;
        uint32_t word = (w < nw) ? mask[w0 + w] : 0u;
        const int cnt = __popc(word);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        if (total == 0) continue;
        unsigned long long pos = 0;
        if (lane == 0) pos = atomicAdd(counter, (unsigned long long)total);
        pos = __shfl_sync(0xffffffffu, pos, 0) + (unsigned long long)(incl - cnt);
        const int32_t row0 = (int32_t)((w0 + w) << 5);
        while (word) {
            const int b = __ffs(word) - 1;
            word &= word - 1u;
            worklist[pos++] = row0 + b;
        }
    }
}

int launch_compact_worklist(kmc_ctx* c, int32_t* worklist, unsigned long long* amb_counter, int64_t r0, int64_t r1,
                            const unsigned int* gate_flag, unsigned int gate_want, cudaStream_t s) {
    const int64_t w0 = r0 >> 5, nw = ((r1 + 31) >> 5) - w0;
    if (nw <= 0) return 0;
    int64_t blocks = (nw + 255) / 256;
    const int64_t cap = (int64_t)c->sm_count * 8;
    blocks = blocks > cap ? cap : blocks;
    compact_worklist_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->at<uint32_t>(c->off_ambmask), w0, nw, worklist, amb_counter,
                                                             gate_flag, gate_want);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

// Device-side decision between the two precisions, per row chunk and pass.  What the uncertified rows of a 1-term
// pass cost depends on their refine walks (measured at config 4, one B200, ns per row at full occupancy): a short
// walk of clustered data (2-3 exact distances) 0.2 ns, a walk up to its cap ~1 ns, a row that overflows the cap
// and goes to the register-tiled exact kernel 21 ns; the 3-term pass costs 0.37 ns more per row of the CHUNK than
// the 1-term pass (61 against 24 ms per 100M rows) -- both of the last two scale with K*D.  So 1 term + refine wins at
// any uncertified fraction while walks end early, and loses once overflowing rows exceed ~1.8% of the chunk.
//   normal pass: the 1-term kernel has run; expected overflow = its uncertified rows x the overflow share of the
//                previous refine (which also followed a 1-term pass).  Above 1.8% of the chunk: the 3-term kernel
//                redoes the chunk, and the next `period` passes go straight to it (period doubles up to 16 while
//                fallbacks recur, and restarts at 1 after a pass that stayed with 1 term).
//   skip pass:   the 1-term kernel did not run; 3 terms.
//   probe pass:  the first pass after a skip period keeps its 1-term result unconditionally, so that its refine
//                measures the overflow share afresh (statistics of 3-term leftovers say nothing about 1-term rows).
__global__ void tc_gate_kernel(DeviceStats* stats, int ch, unsigned long long rows, int prev_slot, int acc16_policy) {
    unsigned long long* amb = &stats->n_amb[ch];
    const bool was16 = acc16_policy && stats->acc16_off[ch] == 0u;      // the tier the 1-term pass of this chunk ran with
    if (stats->tc_skip[ch] > 0u) {
        if (--stats->tc_skip[ch] == 0u) {
            stats->tc_probe[ch] = 1u;
            if (acc16_policy) {          // the probe runs with fp32 accumulators (they certify more: a cheaper probe on
                stats->acc16_off[ch] = 1u;   // hard data); if the pass after it stays with one term, fp16 is tried again
                stats->t1_ok[ch] = 1u;
            }
        }
        stats->tc3_flag[ch] = 1u;
        *amb = 0ull;
        return;
    }
    if (stats->tc_probe[ch]) {
        stats->tc_probe[ch] = 0u;
        stats->tc3_flag[ch] = 0u;
        return;
    }
    const unsigned long long refined = stats->walk[prev_slot][1], over = stats->walk[prev_slot][2];
    const double expect_over = refined > 0ull ? (double)*amb * ((double)over / (double)refined) : 0.0;
    unsigned int period = stats->tc_period[ch] ? stats->tc_period[ch] : 1u;
    if (expect_over > 0.018 * (double)rows) {
        stats->tc3_flag[ch] = 1u;
        *amb = 0ull;
        if (was16) {
            // the fp16 tier left too much to an expensive refine: the fp32-accumulator kernel takes over as first tier
            // (no skip period: it certifies more rows and gets its own chance at the next pass) until `t1_ok` more
            // passes stayed with it; that count (`rearm`) doubles with every failure of the fp16 tier, up to 32, and
            // restarts at 2 when an fp16 pass goes through
            const unsigned int r = stats->rearm[ch] ? stats->rearm[ch] : 2u;
            stats->acc16_off[ch] = 1u;
            stats->t1_ok[ch] = r;
            stats->rearm[ch] = r < 32u ? 2u * r : 32u;
        } else {
            stats->tc_skip[ch] = period;
            stats->tc_period[ch] = period < 16u ? 2u * period : 16u;
        }
    } else {
        stats->tc3_flag[ch] = 0u;
        stats->tc_period[ch] = 1u;
        if (was16) stats->rearm[ch] = 2u;
        if (acc16_policy && !was16) {       // a pass that stayed with the fp32-accumulator tier: towards re-arming fp16
            if (stats->t1_ok[ch] <= 1u) { stats->acc16_off[ch] = 0u; stats->t1_ok[ch] = 0u; }
            else --stats->t1_ok[ch];
        }
    }
}

int launch_tc_gate(kmc_ctx* c, int ch, int64_t rows, bool acc16_policy, cudaStream_t s) {
    tc_gate_kernel<<<1, 1, 0, s>>>(c->stats(), ch, (unsigned long long)rows, (int)((c->refine_launches & 1u) ^ 1u),
                                   acc16_policy ? 1 : 0);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

template <int KC>
static int launch_tc_kc(kmc_ctx* c, const TcParams& p, int terms, cudaStream_t s) {
    const int threads = 128 + 32 * 8;
    const int grid = p.n_mtiles < c->sm_count ? p.n_mtiles : c->sm_count;
    auto launch = [&](auto kern, uint32_t smem_bytes, const TcParams& pp) -> int {
        KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        kern<<<grid, threads, smem_bytes, s>>>(c->tmap_xhi, c->tmap_xlo, c->tmap_c, c->tmap_bias, c->tmap_xext, pp);
        KMC_CUDA(cudaGetLastError());
        c->launches++;
        return 0;
    };
    // CTA pairs: clusters of two CTAs (one per SM of a TPC), the centroid maps with 128-row boxes
    auto launch_pair = [&](auto kern, uint32_t smem_bytes, const TcParams& pp) -> int {
        KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        const int pairs_needed = (p.n_mtiles + 1) / 2, pairs_max = c->sm_count / 2;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(2 * (pairs_needed < pairs_max ? pairs_needed : pairs_max)));
        cfg.blockDim = dim3((unsigned)threads);
        cfg.dynamicSmemBytes = smem_bytes;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (c->tc_tile_n == 128) {
            TcParams p4 = pp;
            p4.n_ntiles = pp.n_ntiles * 2;
            KMC_CUDA(cudaLaunchKernelEx(&cfg, kern, c->tmap_xhi, c->tmap_xlo, c->tmap_c_quarter, c->tmap_bias_quarter, c->tmap_xext, p4));
        } else {
            KMC_CUDA(cudaLaunchKernelEx(&cfg, kern, c->tmap_xhi, c->tmap_xlo, c->tmap_c_half, c->tmap_bias_half, c->tmap_xext, pp));
        }
        c->launches++;
        return 0;
    };
    TcParams pp = p;
    pp.use_bias_mma = 0;
    if (terms == 16) {
        // fp16 accumulators (1 term, bias through the MMA at this tier's own scale): its own centroid / bias maps
        auto launch16 = [&](auto kern, uint32_t smem_bytes) -> int {
            KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
            kern<<<grid, threads, smem_bytes, s>>>(c->tmap_xhi, c->tmap_xlo, c->tmap_c16, c->tmap_bias16, c->tmap_xext, pp);
            KMC_CUDA(cudaGetLastError());
            c->launches++;
            return 0;
        };
        if (p.scores) return launch16(tc_assign_kernel<KC, true, 1, 8, true, 1, 256, true>, TcCfg<KC, 1, 1, 256, true>::kSmemBytes);
        if (c->tc_pair && p.n_mtiles >= 2 && p.n_ntiles * KC >= 4) {
            auto kern = tc_assign_kernel<KC, false, 1, 8, true, 2, 256, true>;
            const uint32_t smem_bytes = TcCfg<KC, 1, 2, 256, true>::kSmemBytes;
            KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
            const int pairs_needed = (p.n_mtiles + 1) / 2, pairs_max = c->sm_count / 2;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(2 * (pairs_needed < pairs_max ? pairs_needed : pairs_max)));
            cfg.blockDim = dim3((unsigned)threads);
            cfg.dynamicSmemBytes = smem_bytes;
            cfg.stream = s;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            KMC_CUDA(cudaLaunchKernelEx(&cfg, kern, c->tmap_xhi, c->tmap_xlo, c->tmap_c16_half, c->tmap_bias16_half, c->tmap_xext, pp));
            c->launches++;
            return 0;
        }
        return launch16(tc_assign_kernel<KC, false, 1, 8, true, 1, 256, true>, TcCfg<KC, 1, 1, 256, true>::kSmemBytes);
    }
    if (p.scores) return launch(tc_assign_kernel<KC, true, 3, 8, false, 1>, TcCfg<KC, 3>::kSmemBytes, pp);
    if (terms != 1) return launch(tc_assign_kernel<KC, false, 3, 8, false, 1>, TcCfg<KC, 3>::kSmemBytes, pp);
#ifdef KMC_NO_BIAS_MMA      // A/B build: biases added by the epilogue (FADD per score) instead of the extra MMA step
    return launch(tc_assign_kernel<KC, false, 1, 8, false, 1>, TcCfg<KC, 1>::kSmemBytes, pp);
#endif
    // one term: the variant that lets the MMA add the biases, and the plain one; the device runs exactly one
    pp.use_bias_mma = 1;
    // pairs pay off when a row tile meets enough MMA work to hide the cross-CTA hand-offs (config 3, one stage per
    // row tile, is faster on single CTAs)
    if (c->tc_pair && p.n_mtiles >= 2 && p.n_ntiles * KC >= 4) {
        if (c->tc_tile_n == 128)
            KMC_TRY(launch_pair(tc_assign_kernel<KC, false, 1, 8, true, 2, 128>, TcCfg<KC, 1, 2, 128>::kSmemBytes, pp));
        else
            KMC_TRY(launch_pair(tc_assign_kernel<KC, false, 1, 8, true, 2, 256>, TcCfg<KC, 1, 2, 256>::kSmemBytes, pp));
    }
    else
        KMC_TRY(launch(tc_assign_kernel<KC, false, 1, 8, true, 1>, TcCfg<KC, 1>::kSmemBytes, pp));
    return launch(tc_assign_kernel<KC, false, 1, 8, false, 1>, TcCfg<KC, 1>::kSmemBytes, pp);
}

int launch_tc_assign(kmc_ctx* c, int32_t* labels_out, int32_t* worklist, unsigned long long* amb_counter,
                     float* scores_out, int64_t r0, int64_t r1, int terms, const unsigned int* gate_flag,
                     unsigned int gate_want, cudaStream_t s, const unsigned int* gate2, unsigned int gate2_want) {
    if (!c->tmaps_ready) KMC_FAIL(KMC_E_STATE, "tensor maps not built (bind workspace and points first)");
    if (r1 <= r0) return 0;
    if (r0 % kTcBlockM) KMC_FAIL(KMC_E_INVALID, "row chunks of the tensor kernel start at multiples of %d", kTcBlockM);
    TcParams p;
    p.n = r1;
    p.tile0 = (int)(r0 / kTcBlockM);
    p.amb_counter = amb_counter;
    p.gate_flag = gate_flag;
    p.gate_want = gate_want;
    p.gate2 = gate2;
    p.gate2_want = gate2_want;
    p.K = (int)c->K;
    p.K_pad = (int)c->K_pad;
    p.n_mtiles = (int)((r1 - r0 + kTcBlockM - 1) / kTcBlockM);
    p.n_ntiles = (int)(c->K_pad / kTcBlockN);
    p.xnorm = c->at<float>(c->off_xnorm);
    p.cnorm = c->at<float>(c->off_cnorm);
    p.stats = c->stats();
    p.labels = labels_out;
    p.worklist = worklist;
    p.amb_mask = c->at<uint32_t>(c->off_ambmask);
    p.scores = scores_out;
    p.gamma = (float)c->tc_gamma;
    p.gamma1 = (float)(1.25 * ((double)(c->D_pad / 16) + 6.0) / 8388608.0);
    // one rounding to fp16 (2^-11 of the accumulator, round to nearest) per MMA instruction of a tile, 25 % head-room;
    // measured on the GPU (tools/check_acc16.py): the largest score error is 0.47 of this bound
    p.gamma16 = (float)(1.25 * (double)(c->D_pad / 16) / 2048.0);
    p.pack_mask = 0xFFFFFF80u;
    p.blk_mask = ~(uint32_t)(kEpiBlocksPerHalf - 1);
    p.x = c->points;
    p.cent = c->centroids();
    p.D = (int)c->D;
    p.use_bias_mma = 0;
    if (!gate_flag) KMC_CUDA(cudaMemsetAsync(amb_counter, 0, sizeof(unsigned long long), s));
    switch ((int)(c->D_pad / kTcChunkK)) {
        case 1: return launch_tc_kc<1>(c, p, terms, s);
        case 2: return launch_tc_kc<2>(c, p, terms, s);
        case 4: return launch_tc_kc<4>(c, p, terms, s);
        case 3: return launch_tc_kc<4>(c, p, terms, s);   // D_pad is rounded to 256 for 3 chunks at create time
        default: KMC_FAIL(KMC_E_UNSUPPORTED, "tensor path supports D <= 256 (D_pad=%lld)", (long long)c->D_pad);
    }
}

}  // namespace kmc

#ifdef KMC_TRACE
extern "C" int kmc_debug_trace(long long* host_out) {      // 256 x 8 stamps of the last tensor-kernel launch
    return (int)cudaMemcpyFromSymbol(host_out, kmc::g_trace, sizeof(long long) * 256 * 8);
}
#endif
