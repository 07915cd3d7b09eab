// Small-D path (HBM-bound configs, e.g. D=2/K=8 and D=3/K=16): ONE pass over the points does
// distance, argmin and accumulation.  Centroids live in shared memory; each thread owns four
// consecutive rows (4*D floats = D aligned float4 loads) and a private column of shared-memory
// accumulators (conflict-free: bank = thread id), folded once per CTA into fp64 global partials.
//
// Replaces reference scripts/distribuitedClustering.py:221-234 (tile/sub/square/sum/argmin on
// materialised [N,K,M] tensors), :237-242 (K x gather+mean) and :244-248 (CPU bincount, mean*count).
// Distances use the reference's exact difference form sum_j (x_j - c_j)^2 (no cancellation);
// ties go to the lowest centroid index (strict '<', ascending k), NaN centroids never win.
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

constexpr int kSmallThreads = 128;
constexpr int kSmallMaxSlots = 384;   // K*(D+1) accumulator slots per thread column

bool small_path_supported(int64_t K, int64_t D) {
    return D >= 1 && D <= 8 && K >= 1 && K * (D + 1) <= kSmallMaxSlots;
}

template <int D, bool ACCUM, bool LABELS>
__global__ void __launch_bounds__(kSmallThreads)
small_kernel(const float* __restrict__ x, int64_t n, const float* __restrict__ cent, int K,
             double* __restrict__ partials, int32_t* __restrict__ labels, int vec_ok) {
    extern __shared__ __align__(16) float smem_f[];
    float* cs = smem_f;                                   // [K*D] centroids
    float* acc = smem_f + ((K * D + 3) & ~3);             // [K*(D+1)][kSmallThreads]
    const int tid = threadIdx.x;
    const int S = K * (D + 1);

    for (int i = tid; i < K * D; i += kSmallThreads) cs[i] = cent[i];
    if (ACCUM) {
        for (int s = 0; s < S; ++s) acc[s * kSmallThreads + tid] = 0.f;
    }
    __syncthreads();

    float inertia = 0.f;
    const int64_t n_groups = (n + 3) >> 2;
    for (int64_t g = (int64_t)blockIdx.x * kSmallThreads + tid; g < n_groups;
         g += (int64_t)gridDim.x * kSmallThreads) {
        const int64_t row0 = g << 2;
        const int valid = (n - row0 >= 4) ? 4 : (int)(n - row0);
        float xs[4 * D];
        if (valid == 4 && vec_ok) {
            const float4* p = reinterpret_cast<const float4*>(x + row0 * D);
#pragma unroll
            for (int q = 0; q < D; ++q) {
                float4 v = ld_stream_f4(p + q);
                xs[4 * q + 0] = v.x; xs[4 * q + 1] = v.y; xs[4 * q + 2] = v.z; xs[4 * q + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 4 * D; ++e)
                xs[e] = (e < valid * D) ? x[row0 * D + e] : 0.f;
        }
        float best[4];
        int bk[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) { best[r] = INFINITY; bk[r] = 0; }
#pragma unroll 4
        for (int k = 0; k < K; ++k) {
            float c[D];
#pragma unroll
            for (int j = 0; j < D; ++j) c[j] = cs[k * D + j];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                float t = xs[r * D] - c[0];
                float d = t * t;
#pragma unroll
                for (int j = 1; j < D; ++j) {
                    t = xs[r * D + j] - c[j];
                    d = fmaf(t, t, d);
                }
                if (d < best[r]) { best[r] = d; bk[r] = k; }
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
                    float* a = acc + (bk[r] * (D + 1)) * kSmallThreads + tid;
#pragma unroll
                    for (int j = 0; j < D; ++j) a[j * kSmallThreads] += xs[r * D + j];
                    a[D * kSmallThreads] += 1.f;
                    inertia += best[r];
                }
            }
        }
    }

    if (ACCUM) {
        __syncthreads();
        const int lane = tid & 31, warp = tid >> 5;
        for (int s = warp; s < S; s += kSmallThreads / 32) {
            double v = 0.0;
#pragma unroll
            for (int q = 0; q < kSmallThreads / 32; ++q) v += (double)acc[s * kSmallThreads + q * 32 + lane];
            v = warp_sum(v);
            if (lane == 0 && v != 0.0) {
                const int k = s / (D + 1), j = s - k * (D + 1);
                double* dst = (j < D) ? (partials + (int64_t)k * D + j) : (partials + (int64_t)K * D + k);
                atomicAdd(dst, v);
            }
        }
        double in = warp_sum((double)inertia);
        if (lane == 0 && in != 0.0) atomicAdd(partials + (int64_t)K * D + K, in);
    }
}

template <int D>
static int launch_small_d(kmc_ctx* c, bool accumulate, int32_t* labels_out, cudaStream_t s) {
    const int K = (int)c->K;
    const size_t smem_c = (size_t)((K * D + 3) & ~3) * sizeof(float);
    const size_t smem_acc = (size_t)K * (D + 1) * kSmallThreads * sizeof(float);
    const size_t smem = smem_c + (accumulate ? smem_acc : 0);
    const int64_t n_groups = (c->n + 3) / 4;
    int64_t blocks = (n_groups + kSmallThreads - 1) / kSmallThreads;
    int per_sm = (int)((200 * 1024) / (smem + 1024));
    per_sm = per_sm < 1 ? 1 : (per_sm > 12 ? 12 : per_sm);
    const int64_t cap = (int64_t)c->sm_count * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(c->points) & 15) == 0) &&
                       (labels_out == nullptr || (reinterpret_cast<uintptr_t>(labels_out) & 15) == 0);
    auto launch = [&](auto kern) -> int {
        KMC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)blocks, kSmallThreads, smem, s>>>(c->points, c->n, c->centroids(), K, c->partials(),
                                                           labels_out, vec_ok);
        KMC_CUDA(cudaGetLastError());
        c->launches++;
        return 0;
    };
    if (accumulate && labels_out) return launch(small_kernel<D, true, true>);
    if (accumulate) return launch(small_kernel<D, true, false>);
    return launch(small_kernel<D, false, true>);
}

int launch_small(kmc_ctx* c, bool accumulate, int32_t* labels_out, cudaStream_t s) {
    if (!accumulate && !labels_out) return 0;
    switch ((int)c->D) {
        case 1: return launch_small_d<1>(c, accumulate, labels_out, s);
        case 2: return launch_small_d<2>(c, accumulate, labels_out, s);
        case 3: return launch_small_d<3>(c, accumulate, labels_out, s);
        case 4: return launch_small_d<4>(c, accumulate, labels_out, s);
        case 5: return launch_small_d<5>(c, accumulate, labels_out, s);
        case 6: return launch_small_d<6>(c, accumulate, labels_out, s);
        case 7: return launch_small_d<7>(c, accumulate, labels_out, s);
        case 8: return launch_small_d<8>(c, accumulate, labels_out, s);
        default: KMC_FAIL(KMC_E_UNSUPPORTED, "small path needs 1 <= D <= 8 (got %lld)", (long long)c->D);
    }
}

}  // namespace kmc
