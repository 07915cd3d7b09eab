// Small-D path (HBM-bound configs, e.g. D=2/K=8 and D=3/K=16): ONE pass over the points does
// distance, argmin and accumulation.  Centroids live in shared memory; each thread owns four
// consecutive rows (4*D floats = D aligned float4 loads, the next group prefetched while the current one
// is processed) and a private column of shared-memory accumulators (conflict-free: bank = thread id),
// folded once per CTA into fp64 global partials.  The grid is persistent (two CTAs per SM), so the
// number of fp64 atomics per partial is a few hundred, not one per tile of rows.
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

constexpr int kSmallThreads = 256;
constexpr int kSmallMaxSlots = 384;   // K*(D+1) accumulator slots per thread column (at 128 threads)

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

// FUSE (kmc_iterate on one GPU): the last CTA to flush its sums also does the centroid update (reference
// :257-263) and clears the partial buffer of the NEXT iteration, so a whole Lloyd iteration is ONE launch and no
// memset -- at config 1 (1M rows, 8 MB, L2-resident) launches and memsets were half of the iteration.
struct SmallFuse {
    float* cent_w;            // the centroids again (written by the last CTA; every CTA has read them by then)
    float* cent_prev;
    double* next_partials;    // cleared for the next iteration
    DeviceStats* stats;
    int empty_nan;
};

template <int D, int THREADS, bool ACCUM, bool LABELS, bool PACK, bool FUSE>
__global__ void __launch_bounds__(THREADS)
small_kernel(const float* __restrict__ x, int64_t n, const float* __restrict__ cent, int K,
             double* __restrict__ partials, int32_t* __restrict__ labels, int vec_ok, uint32_t idx_mask,
             const SmallFuse fuse) {
    extern __shared__ __align__(16) float smem_f[];
    float* ncs = smem_f;                                  // [K*D] NEGATED centroids
    float* acc = smem_f + ((K * D + 3) & ~3);             // [K*(D+1)][THREADS]
    const int tid = threadIdx.x;
    const int S = K * (D + 1);

    for (int i = tid; i < K * D; i += THREADS) ncs[i] = -cent[i];
    if (ACCUM) {
        for (int s = 0; s < S; ++s) acc[s * THREADS + tid] = 0.f;
    }
    __syncthreads();

    float inertia = 0.f;
    const int64_t n_groups = (n + 3) >> 2;
    const int64_t stride = (int64_t)gridDim.x * THREADS;
    int64_t g = (int64_t)blockIdx.x * THREADS + tid;
    float nxt[4 * D];
    if (g < n_groups) load_group<D>(x, n, g, vec_ok, nxt);
    for (; g < n_groups; g += stride) {
        float f[4 * D];
#pragma unroll
        for (int e = 0; e < 4 * D; ++e) f[e] = nxt[e];
        if (g + stride < n_groups) load_group<D>(x, n, g + stride, vec_ok, nxt);   // in flight during the math
        const int64_t row0 = g << 2;
        const int valid = (n - row0 >= 4) ? 4 : (int)(n - row0);
        // rows (0,1) and (2,3) as packed pairs per dimension
        float2 xp[2][D];
#pragma unroll
        for (int j = 0; j < D; ++j) {
            xp[0][j] = make_float2(f[j], f[D + j]);
            xp[1][j] = make_float2(f[2 * D + j], f[3 * D + j]);
        }
        float best[4];
        int bk[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) { best[r] = INFINITY; bk[r] = 0; }

        auto dist4 = [&](int k, float (&d)[4]) {
            float2 d0, d1;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float nc = ncs[k * D + j];
                const float2 nc2 = make_float2(nc, nc);
                const float2 t0 = __fadd2_rn(xp[0][j], nc2), t1 = __fadd2_rn(xp[1][j], nc2);
                if (j == 0) { d0 = __fmul2_rn(t0, t0); d1 = __fmul2_rn(t1, t1); }
                else { d0 = __ffma2_rn(t0, t0, d0); d1 = __ffma2_rn(t1, t1, d1); }
            }
            d[0] = d0.x; d[1] = d0.y; d[2] = d1.x; d[3] = d1.y;
        };

        if (PACK) {
            int k = 0;
            for (; k + 1 < K; k += 2) {
                float da[4], db[4];
                dist4(k, da);
                dist4(k + 1, db);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float pa = __uint_as_float((__float_as_uint(da[r]) & idx_mask) | (uint32_t)k);
                    const float pb = __uint_as_float((__float_as_uint(db[r]) & idx_mask) | (uint32_t)(k + 1));
                    best[r] = fminf(fminf(best[r], pa), pb);
                }
            }
            if (k < K) {
                float da[4];
                dist4(k, da);
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    best[r] = fminf(best[r], __uint_as_float((__float_as_uint(da[r]) & idx_mask) | (uint32_t)k));
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const uint32_t b = __float_as_uint(best[r]);
                bk[r] = (best[r] < INFINITY) ? (int)(b & ~idx_mask) : 0;
                best[r] = __uint_as_float(b & idx_mask);
            }
        } else {
#pragma unroll 2
            for (int k = 0; k < K; ++k) {
                float d[4];
                dist4(k, d);
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (d[r] < best[r]) { best[r] = d[r]; bk[r] = k; }
            }
        }
        if (LABELS) {
            if (valid == 4 && vec_ok) {
                *reinterpret_cast<int4*>(labels + row0) = make_int4(bk[0], bk[1], bk[2], bk[3]);
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (r < valid) labels[row0 + r] = bk[r];
            }
        }
        if (ACCUM) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (r < valid) {
                    float* a = acc + (bk[r] * (D + 1)) * THREADS + tid;
#pragma unroll
                    for (int j = 0; j < D; ++j) a[j * THREADS] += f[r * D + j];
                    a[D * THREADS] += 1.f;
                    if (best[r] < INFINITY) inertia += best[r];
                }
            }
        }
    }

    if (ACCUM) {
        __syncthreads();
        const int lane = tid & 31, warp = tid >> 5;
        for (int s = warp; s < S; s += THREADS / 32) {
            double v = 0.0;
#pragma unroll
            for (int q = 0; q < THREADS / 32; ++q) v += (double)acc[s * THREADS + q * 32 + lane];
            v = warp_sum(v);
            if (lane == 0 && v != 0.0) {
                const int k = s / (D + 1), j = s - k * (D + 1);
                double* dst = (j < D) ? (partials + (int64_t)k * D + j) : (partials + (int64_t)K * D + k);
                atomicAdd(dst, v);
            }
        }
        double in = warp_sum((double)inertia);
        if (lane == 0 && in != 0.0) atomicAdd(partials + (int64_t)K * D + K, in);
        if (FUSE) {
            __shared__ unsigned int s_last;
            __shared__ double s_shift[THREADS / 32];
            __threadfence();
            __syncthreads();
            if (tid == 0) s_last = (atomicAdd(&fuse.stats->upd_done, 1u) == gridDim.x - 1) ? 1u : 0u;
            __syncthreads();
            if (s_last) {
                __threadfence();
                const int KD = K * D;
                double shift = 0.0;
                for (int i = tid; i < KD; i += THREADS) {
                    const double cnt = __ldcg(partials + KD + i / D);
                    const float old = -ncs[i];
                    fuse.cent_prev[i] = old;
                    const float nw = (cnt > 0.0) ? (float)(__ldcg(partials + i) / cnt)
                                                 : (fuse.empty_nan ? __int_as_float(0x7fc00000) : old);
                    fuse.cent_w[i] = nw;
                    const double d = (double)nw - (double)old;
                    if (d == d) shift += d * d;
                }
                shift = warp_sum(shift);
                if (lane == 0) s_shift[warp] = shift;
                for (int i = tid; i < KD + K + 2; i += THREADS) fuse.next_partials[i] = 0.0;
                __syncthreads();
                if (tid == 0) {
                    double t = 0.0;
                    for (int w = 0; w < THREADS / 32; ++w) t += s_shift[w];
                    fuse.stats->shift2 = t;
                    fuse.stats->upd_done = 0u;
                }
            }
        }
    }
}

template <int D, int THREADS>
static int launch_small_dt(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                           double* partials, double* fuse_next) {
    const int K = (int)c->K;
    const size_t smem_c = (size_t)((K * D + 3) & ~3) * sizeof(float);
    const size_t smem_acc = (size_t)K * (D + 1) * THREADS * sizeof(float);
    const size_t smem = smem_c + (accumulate ? smem_acc : 0);
    const int64_t n_groups = (c->n + 3) / 4;
    int64_t blocks = (n_groups + THREADS - 1) / THREADS;
    // persistent grid: as many CTAs as stay resident, up to 1024 threads per SM (the loads in flight per SM,
    // one 16*D-byte group per thread, are what reaches HBM bandwidth)
    int per_sm = (int)((200 * 1024) / (smem + 1024));
    const int want = 1024 / THREADS;
    per_sm = per_sm < 1 ? 1 : (per_sm > want ? want : per_sm);
    const int64_t cap = (int64_t)c->sm_count * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(c->points) & 15) == 0) &&
                       (labels_out == nullptr || (reinterpret_cast<uintptr_t>(labels_out) & 15) == 0);
    int bits = 1;
    while ((1 << bits) < K) ++bits;
    const bool pack = K <= 64;
    const uint32_t idx_mask = ~((1u << bits) - 1u);
    SmallFuse fuse = {c->centroids(), c->centroids_prev(), fuse_next, c->stats(), (c->flags & KMC_EMPTY_NAN) ? 1 : 0};
    auto launch = [&](auto kern) -> int {
        KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)blocks, THREADS, smem, s>>>(c->points, c->n, cent, K, partials, labels_out,
                                                     vec_ok, idx_mask, fuse);
        KMC_CUDA(cudaGetLastError());
        c->launches++;
        return 0;
    };
    if (fuse_next) {      // one launch per Lloyd iteration: accumulate + update (no labels)
        if (pack) return launch(small_kernel<D, THREADS, true, false, true, true>);
        return launch(small_kernel<D, THREADS, true, false, false, true>);
    }
    if (pack) {
        if (accumulate && labels_out) return launch(small_kernel<D, THREADS, true, true, true, false>);
        if (accumulate) return launch(small_kernel<D, THREADS, true, false, true, false>);
        return launch(small_kernel<D, THREADS, false, true, true, false>);
    }
    if (accumulate && labels_out) return launch(small_kernel<D, THREADS, true, true, false, false>);
    if (accumulate) return launch(small_kernel<D, THREADS, true, false, false, false>);
    return launch(small_kernel<D, THREADS, false, true, false, false>);
}

template <int D>
static int launch_small_d(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                          double* partials, double* fuse_next) {
    // 256 threads per CTA when the private accumulator columns fit twice per SM, else 128
    const size_t acc256 = (size_t)c->K * (D + 1) * 256 * sizeof(float);
    if (!accumulate || acc256 <= 96 * 1024) return launch_small_dt<D, kSmallThreads>(c, accumulate, labels_out, cent, s, partials, fuse_next);
    return launch_small_dt<D, 128>(c, accumulate, labels_out, cent, s, partials, fuse_next);
}

int launch_small(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s) {
    return launch_small_into(c, accumulate, labels_out, cent, s, c->partials(), nullptr);
}

// `partials`: where the sums go; `fuse_next` non-null: fused iteration (see SmallFuse), that buffer is cleared
int launch_small_into(kmc_ctx* c, bool accumulate, int32_t* labels_out, const float* cent, cudaStream_t s,
                      double* partials, double* fuse_next) {
    if (!accumulate && !labels_out) return 0;
    switch ((int)c->D) {
        case 1: return launch_small_d<1>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 2: return launch_small_d<2>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 3: return launch_small_d<3>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 4: return launch_small_d<4>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 5: return launch_small_d<5>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 6: return launch_small_d<6>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 7: return launch_small_d<7>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        case 8: return launch_small_d<8>(c, accumulate, labels_out, cent, s, partials, fuse_next);
        default: KMC_FAIL(KMC_E_UNSUPPORTED, "small path needs 1 <= D <= 8 (got %lld)", (long long)c->D);
    }
}

}  // namespace kmc
