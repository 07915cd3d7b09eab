// Small-D path (HBM-bound configs, e.g. D=2/K=8 and D=3/K=16): ONE pass over the points does
// distance, argmin and accumulation.  Centroids live in shared memory; each thread owns four
// consecutive rows (4*D floats = D aligned float4 loads, the next group prefetched while the current one
// is processed) and a private column of shared-memory accumulators (conflict-free: bank = thread id),
// folded once per CTA into fp64.
//
// Grid: ONE CTA per SM, as many threads (128 ... 1024) as private accumulator columns fit into shared memory.
// No atomics and no fences touch the sums: every CTA stores its K*(D+1)+1 partial values in its own column of
// a slot table in the workspace (148 columns), and a one-CTA fold kernel launched right behind adds the columns up
// in a fixed order (bit-reproducible) and, for kmc_iterate, updates the centroids.  Both kernels are launched with
// programmatic stream serialisation: a kernel's CTAs are resident and have cleared their accumulators by the time
// the kernel before it drains (griddepcontrol.wait stands between that prologue and the first dependent read).
// Measured at config 1 (1M rows) on the way here: a few hundred CTAs adding into the same 25 doubles with fp64
// atomics plus a __threadfence() per thread for a last-CTA-done hand-off (MEMBAR.SC + ERRBAR + L1 invalidate) were
// 59 % of the kernel's stall samples; with one fence per CTA the last CTA's serial fold of a 592-column table still
// left the SMs idle for two thirds of the kernel's duration.  The kernel boundary is the cheaper synchronisation.
//
// Arithmetic: the reference's exact difference form sum_j (x_j - c_j)^2, evaluated two rows at a time
// with the packed fp32 pipe (add.f32x2 / fma.f32x2).  For K <= 64 the argmin runs on values that carry
// the centroid index in their lowest mantissa bits (one LOP3 + half a 3-input FMNMX per candidate
// instead of compare + two selects); the truncation is < 2^-17 relative, below the 1e-5 near-tie window,
// and equal truncated values resolve to the lowest index.  NaN centroids never win.
//
// Replaces reference scripts/distribuitedClustering.py:221-234 (tile/sub/square/sum/argmin on
// materialised [N,K,M] tensors), :237-242 (K x gather+mean) and :244-248 (CPU bincount, mean*count).
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

constexpr int kSmallMaxSlots = 384;   // K*(D+1) accumulator slots per thread column (at 128 threads)
constexpr size_t kSmallSmemBudget = 200 * 1024;

// CTAs of the persistent grid at most (= columns of the slot table in the workspace): one per SM
int64_t small_max_ctas(int sm_count) { return (int64_t)sm_count; }

// Programmatic dependent launch: wait until the kernels this launch depends on have completed and their writes
// are visible / let the next kernel in the stream start its prologue.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool small_path_supported(int64_t K, int64_t D) {
    return D >= 1 && D <= 8 && K >= 1 && K * (D + 1) <= kSmallMaxSlots;
}

template <int D>
__device__ __forceinline__ void load_group(const float* __restrict__ x, int64_t n, int64_t g, int vec_ok,
                                           float (&f)[4 * D]) {
    const int64_t row0 = g << 2;
    if (n - row0 >= 4 && vec_ok) {
        const float4* p = reinterpret_cast<const float4*>(x + row0 * D);
#pragma unroll
        for (int q = 0; q < D; ++q) {
            const float4 v = ld_stream_f4(p + q);
            f[4 * q + 0] = v.x; f[4 * q + 1] = v.y; f[4 * q + 2] = v.z; f[4 * q + 3] = v.w;
        }
    } else {
        const int valid = (n - row0 >= 4) ? 4 : (int)(n - row0);
#pragma unroll
        for (int e = 0; e < 4 * D; ++e) f[e] = (e < valid * D) ? x[row0 * D + e] : 0.f;
    }
}

// What the fold kernel needs besides the slot table.  FUSE (kmc_iterate on one GPU): it also does the centroid
// update (reference :257-263), so a Lloyd iteration is two launches, no memset and no host round trip.  Without
// FUSE it ADDS the totals to the partial buffer (kmc_accumulate clears it first, kmc_accumulate_add does not).
struct SmallFuse {
    float* cent_w;            // centroids, rewritten by the fused update
    float* cent_prev;
    DeviceStats* stats;
    int empty_nan;
};

// MAXT: upper bound of the block size this instance is launched with (1024, 768 or 512 threads: 64, 80 or 128
// registers per thread)
template <int D, bool ACCUM, bool LABELS, bool PACK, int MAXT>
__global__ void __launch_bounds__(MAXT)
small_kernel(const float* __restrict__ x, int64_t n, const float* __restrict__ cent, int K,
             double* __restrict__ slots, int32_t* __restrict__ labels, int vec_ok, uint32_t idx_mask) {
    extern __shared__ __align__(16) float smem_f[];
    float* ncs = smem_f;                                  // [K*D] NEGATED centroids
    float* acc = smem_f + ((K * D + 3) & ~3);             // [K*(D+1)][THREADS]
    const int tid = threadIdx.x;
    const int THREADS = blockDim.x;
    const int S = K * (D + 1);

    pdl_launch_dependents();          // the fold kernel may become resident; it waits for this grid to complete
    if (ACCUM) {
        for (int s = 0; s < S; ++s) acc[s * THREADS + tid] = 0.f;
    }
    pdl_wait();                       // centroids (fused update of the previous iteration) and the slot table are ours now
    for (int i = tid; i < K * D; i += THREADS) ncs[i] = -__ldcg(cent + i);
    __syncthreads();

    // NG groups of four rows per thread and loop iteration (they would share every centroid load).  Measured on one
    // B200: NG = 2 is SLOWER (config 2: 61.1 us against 56.2 us per pass, config 1: 10.3 against 9.5) -- twice the
    // registers per thread for half the threads' worth of loads in flight; the code stays general, NG stays 1.
    constexpr int NG = 1;
    float inertia = 0.f;
    const int64_t n_groups = (n + 3) >> 2;
    const int64_t stride = (int64_t)gridDim.x * THREADS;
    int64_t g = (int64_t)blockIdx.x * THREADS + tid;
    float nxt[NG][4 * D];
#pragma unroll
    for (int u = 0; u < NG; ++u)
        if (g + u * stride < n_groups) load_group<D>(x, n, g + u * stride, vec_ok, nxt[u]);
    for (; g < n_groups; g += NG * stride) {
        float f[NG][4 * D];
#pragma unroll
        for (int u = 0; u < NG; ++u)
#pragma unroll
            for (int e = 0; e < 4 * D; ++e) f[u][e] = nxt[u][e];
#pragma unroll
        for (int u = 0; u < NG; ++u)          // in flight during the math
            if (g + (NG + u) * stride < n_groups) load_group<D>(x, n, g + (NG + u) * stride, vec_ok, nxt[u]);
        // rows (0,1) and (2,3) of every group as packed pairs per dimension
        float2 xp[NG][2][D];
#pragma unroll
        for (int u = 0; u < NG; ++u)
#pragma unroll
            for (int j = 0; j < D; ++j) {
                xp[u][0][j] = make_float2(f[u][j], f[u][D + j]);
                xp[u][1][j] = make_float2(f[u][2 * D + j], f[u][3 * D + j]);
            }
        float best[NG][4];
        int bk[NG][4];
#pragma unroll
        for (int u = 0; u < NG; ++u)
#pragma unroll
            for (int r = 0; r < 4; ++r) { best[u][r] = INFINITY; bk[u][r] = 0; }

        auto dist4 = [&](int k, float (&d)[NG][4]) {
            float2 d0[NG], d1[NG];
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float nc = ncs[k * D + j];
                const float2 nc2 = make_float2(nc, nc);
#pragma unroll
                for (int u = 0; u < NG; ++u) {
                    const float2 t0 = __fadd2_rn(xp[u][0][j], nc2), t1 = __fadd2_rn(xp[u][1][j], nc2);
                    if (j == 0) { d0[u] = __fmul2_rn(t0, t0); d1[u] = __fmul2_rn(t1, t1); }
                    else { d0[u] = __ffma2_rn(t0, t0, d0[u]); d1[u] = __ffma2_rn(t1, t1, d1[u]); }
                }
            }
#pragma unroll
            for (int u = 0; u < NG; ++u) { d[u][0] = d0[u].x; d[u][1] = d0[u].y; d[u][2] = d1[u].x; d[u][3] = d1[u].y; }
        };

        if (PACK) {
            int k = 0;
            for (; k + 1 < K; k += 2) {
                float da[NG][4], db[NG][4];
                dist4(k, da);
                dist4(k + 1, db);
#pragma unroll
                for (int u = 0; u < NG; ++u)
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const float pa = __uint_as_float((__float_as_uint(da[u][r]) & idx_mask) | (uint32_t)k);
                        const float pb = __uint_as_float((__float_as_uint(db[u][r]) & idx_mask) | (uint32_t)(k + 1));
                        best[u][r] = fminf(fminf(best[u][r], pa), pb);
                    }
            }
            if (k < K) {
                float da[NG][4];
                dist4(k, da);
#pragma unroll
                for (int u = 0; u < NG; ++u)
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        best[u][r] = fminf(best[u][r], __uint_as_float((__float_as_uint(da[u][r]) & idx_mask) | (uint32_t)k));
            }
#pragma unroll
            for (int u = 0; u < NG; ++u)
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t b = __float_as_uint(best[u][r]);
                    bk[u][r] = (best[u][r] < INFINITY) ? (int)(b & ~idx_mask) : 0;
                    best[u][r] = __uint_as_float(b & idx_mask);
                }
        } else {
#pragma unroll 2
            for (int k = 0; k < K; ++k) {
                float d[NG][4];
                dist4(k, d);
#pragma unroll
                for (int u = 0; u < NG; ++u)
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        if (d[u][r] < best[u][r]) { best[u][r] = d[u][r]; bk[u][r] = k; }
            }
        }
#pragma unroll
        for (int u = 0; u < NG; ++u) {
            const int64_t gu = g + u * stride;
            if (gu >= n_groups) break;
            const int64_t row0 = gu << 2;
            const int valid = (n - row0 >= 4) ? 4 : (int)(n - row0);
            if (LABELS) {
                if (valid == 4 && vec_ok) {
                    *reinterpret_cast<int4*>(labels + row0) = make_int4(bk[u][0], bk[u][1], bk[u][2], bk[u][3]);
                } else {
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        if (r < valid) labels[row0 + r] = bk[u][r];
                }
            }
            if (ACCUM) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (r < valid) {
                        float* a = acc + (bk[u][r] * (D + 1)) * THREADS + tid;
#pragma unroll
                        for (int j = 0; j < D; ++j) a[j * THREADS] += f[u][r * D + j];
                        a[D * THREADS] += 1.f;
                        if (best[u][r] < INFINITY) inertia += best[u][r];
                    }
                }
            }
        }
    }

    if (ACCUM) {
        __shared__ double s_red[32];
        __syncthreads();
        const int lane = tid & 31, warp = tid >> 5, nwarps = THREADS >> 5;
        const int64_t G = gridDim.x;
        // this CTA's column of the slot table: slot s = cluster s / (D+1), component s % (D+1) (D = the count)
        for (int s = warp; s < S; s += nwarps) {
            double v = 0.0;
            for (int q = 0; q < nwarps; ++q) v += (double)acc[s * THREADS + q * 32 + lane];
            v = warp_sum(v);
            if (lane == 0) slots[s * G + blockIdx.x] = v;
        }
        const double in_w = warp_sum((double)inertia);
        if (lane == 0) s_red[warp] = in_w;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < nwarps; ++w) t += s_red[w];
            slots[S * G + blockIdx.x] = t;
        }
    }
}

// One CTA of 1024 threads: warp w adds up the rows w, w + 32, ... of the slot table (row = one of the K*(D+1) sums /
// counts or the inertia, column = CTA of small_kernel).  A lane requests all its values of up to three rows before
// it adds any of them (one L2 round trip for config 2's 67 rows x 148 columns), in an order that depends on the
// number of columns only.  FUSE: then the centroid update and the squared shift.
constexpr int kFoldThreads = 1024;
constexpr int kFoldCols = 8;          // columns per lane: the table has at most 32 * kFoldCols = 256 columns
template <bool FUSE>
__global__ void __launch_bounds__(kFoldThreads)
small_fold_kernel(const double* __restrict__ slots, int G, int K, int D, double* __restrict__ partials,
                  const SmallFuse fuse) {
    extern __shared__ double tot[];                   // [K*(D+1)+1] totals, then 32 shift partials
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = K * (D + 1), KD = K * D;
    constexpr int NW = kFoldThreads / 32;
    pdl_launch_dependents();       // the next iteration's small_kernel may start clearing its accumulators
    pdl_wait();                    // every column of the table has been written
    for (int s0 = warp; s0 <= S; s0 += 3 * NW) {
        double v[3][kFoldCols];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int s = s0 + r * NW;
            const double* row = slots + (int64_t)s * G;
#pragma unroll
            for (int u = 0; u < kFoldCols; ++u) {
                const int b = 32 * u + lane;
                v[r][u] = (s <= S && b < G) ? __ldcg(row + b) : 0.0;
            }
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int s = s0 + r * NW;
            double t = ((v[r][0] + v[r][1]) + (v[r][2] + v[r][3])) + ((v[r][4] + v[r][5]) + (v[r][6] + v[r][7]));
            t = warp_sum(t);
            if (lane == 0 && s <= S) {
                tot[s] = t;
                const int k = s / (D + 1), j = s - k * (D + 1);
                double* dst = (s == S) ? (partials + KD + K) : ((j < D) ? (partials + (int64_t)k * D + j) : (partials + KD + k));
                *dst = FUSE ? t : (*dst + t);
            }
        }
    }
    if (!FUSE) return;
    __syncthreads();
    double shift = 0.0;
    for (int i = tid; i < KD; i += kFoldThreads) {
        const int k = i / D, j = i - k * D;
        const double cnt = tot[k * (D + 1) + D];
        const float old = __ldcg(fuse.cent_w + i);
        fuse.cent_prev[i] = old;
        const float nw = (cnt > 0.0) ? (float)(tot[k * (D + 1) + j] / cnt)
                                     : (fuse.empty_nan ? __int_as_float(0x7fc00000) : old);
        fuse.cent_w[i] = nw;
        const double d = (double)nw - (double)old;
        if (d == d) shift += d * d;
    }
    shift = warp_sum(shift);
    double* s_red = tot + S + 1;
    if (lane == 0) s_red[warp] = shift;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < NW; ++w) t += s_red[w];
        fuse.stats->shift2 = t;
    }
}

// launch with programmatic stream serialisation (the kernel calls griddepcontrol.wait before its first dependent access)
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <int D>
static int launch_small_d(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                          double* partials, bool fused) {
    const int K = (int)c->K;
    const size_t smem_c = (size_t)((K * D + 3) & ~3) * sizeof(float);
    // one CTA per SM with as many threads as private accumulator columns (K*(D+1) floats each) fit
    int threads = 1024;
    if (accumulate)
        while (threads > 128 && smem_c + (size_t)K * (D + 1) * threads * sizeof(float) > kSmallSmemBudget) threads -= 128;
    const size_t smem = smem_c + (accumulate ? (size_t)K * (D + 1) * threads * sizeof(float) : 0);
    const int64_t n_groups = (c->n + 3) / 4;
    int64_t blocks = (n_groups + threads - 1) / threads;
    const int64_t cap = small_max_ctas(c->sm_count);         // columns of the slot table
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(c->points) & 15) == 0) &&
                       (labels_out == nullptr || (reinterpret_cast<uintptr_t>(labels_out) & 15) == 0);
    int bits = 1;
    while ((1 << bits) < K) ++bits;
    const bool pack = K <= 64;
    const uint32_t idx_mask = ~((1u << bits) - 1u);
    double* slots = c->at<double>(c->off_small_slots);
    auto launch = [&](auto kern) -> int {
        KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        KMC_CUDA(launch_pdl(kern, (unsigned)blocks, (unsigned)threads, smem, s, c->points, c->n, cent, K, slots, labels_out,
                            vec_ok, idx_mask));
        c->launches++;
        return 0;
    };
#define KMC_SMALL_MODES(MAXT)                                                                         \
    do {                                                                                              \
        if (pack) {                                                                                   \
            if (accumulate && labels_out) KMC_TRY(launch(small_kernel<D, true, true, true, MAXT>));   \
            else if (accumulate) KMC_TRY(launch(small_kernel<D, true, false, true, MAXT>));           \
            else KMC_TRY(launch(small_kernel<D, false, true, true, MAXT>));                           \
        } else {                                                                                      \
            if (accumulate && labels_out) KMC_TRY(launch(small_kernel<D, true, true, false, MAXT>));  \
            else if (accumulate) KMC_TRY(launch(small_kernel<D, true, false, false, MAXT>));          \
            else KMC_TRY(launch(small_kernel<D, false, true, false, MAXT>));                          \
        }                                                                                             \
    } while (0)
    if (threads > 768) KMC_SMALL_MODES(1024);
    else if (threads > 512) KMC_SMALL_MODES(768);
    else KMC_SMALL_MODES(512);
#undef KMC_SMALL_MODES
    if (!accumulate) return 0;
    // fold the slot table into the partial buffer (fused: and update the centroids)
    if (blocks > 32 * kFoldCols) KMC_FAIL(KMC_E_UNSUPPORTED, "slot table of %lld columns (at most %d)", (long long)blocks, 32 * kFoldCols);
    const SmallFuse fuse = {c->centroids(), c->centroids_prev(), c->stats(), (c->flags & KMC_EMPTY_NAN) ? 1 : 0};
    const size_t fold_smem = (size_t)(K * (D + 1) + 1 + kFoldThreads / 32) * sizeof(double);
    if (fused) KMC_CUDA(launch_pdl(small_fold_kernel<true>, 1u, (unsigned)kFoldThreads, fold_smem, s, slots, (int)blocks, K, D, partials, fuse));
    else KMC_CUDA(launch_pdl(small_fold_kernel<false>, 1u, (unsigned)kFoldThreads, fold_smem, s, slots, (int)blocks, K, D, partials, fuse));
    c->launches++;
    return 0;
}

int launch_small(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s) {
    return launch_small_into(c, accumulate, labels_out, cent, s, c->partials(), false);
}

// `partials`: where the sums go; `fused`: the fold kernel overwrites `partials` with the totals and updates the centroids
int launch_small_into(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                      double* partials, bool fused) {
    if (!accumulate && !labels_out) return 0;
    switch ((int)c->D) {
        case 1: return launch_small_d<1>(c, accumulate, labels_out, cent, s, partials, fused);
        case 2: return launch_small_d<2>(c, accumulate, labels_out, cent, s, partials, fused);
        case 3: return launch_small_d<3>(c, accumulate, labels_out, cent, s, partials, fused);
        case 4: return launch_small_d<4>(c, accumulate, labels_out, cent, s, partials, fused);
        case 5: return launch_small_d<5>(c, accumulate, labels_out, cent, s, partials, fused);
        case 6: return launch_small_d<6>(c, accumulate, labels_out, cent, s, partials, fused);
        case 7: return launch_small_d<7>(c, accumulate, labels_out, cent, s, partials, fused);
        case 8: return launch_small_d<8>(c, accumulate, labels_out, cent, s, partials, fused);
        default: KMC_FAIL(KMC_E_UNSUPPORTED, "small path needs 1 <= D <= 8 (got %lld)", (long long)c->D);
    }
}

}  // namespace kmc
