// Staging-side kernels: fp64->fp32 conversion of staged chunks, abs-max scan and the one-time
// split of the resident fp32 points into the two scaled fp16 planes the tcgen05 kernel reads.
//
// Reference: the data lands on the GPU through `X = tf.Variable(X_mat)` + initializer run
// (scripts/distribuitedClustering.py:217, :273); there is no counterpart of the split planes.
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

__global__ void __launch_bounds__(256) convert_f64_kernel(const double* __restrict__ src,
                                                          float* __restrict__ dst, int64_t count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) dst[i] = (float)src[i];
}

int launch_convert_f64(kmc_ctx* c, const double* src, float* dst, int64_t count, cudaStream_t s) {
    if (count <= 0) return 0;
    int64_t blocks = (count + 255) / 256;
    if (blocks > (int64_t)c->sm_count * 16) blocks = (int64_t)c->sm_count * 16;
    convert_f64_kernel<<<(unsigned)blocks, 256, 0, s>>>(src, dst, count);
    KMC_CUDA(cudaGetLastError());
    return 0;
}

// int32 labels -> int64 (the dtype of the reference's np.argmin result), on the device so that the host
// only copies
__global__ void __launch_bounds__(256) widen_labels_kernel(const int32_t* __restrict__ src, long long* __restrict__ dst,
                                                           int64_t count) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) dst[i] = (long long)src[i];
}

int launch_widen_labels(kmc_ctx* c, const int32_t* src, long long* dst, int64_t count, cudaStream_t s) {
    if (count <= 0) return 0;
    int64_t blocks = (count + 255) / 256;
    if (blocks > (int64_t)c->sm_count * 16) blocks = (int64_t)c->sm_count * 16;
    widen_labels_kernel<<<(unsigned)blocks, 256, 0, s>>>(src, dst, count);
    KMC_CUDA(cudaGetLastError());
    return 0;
}

// out[i, :] = centroids[labels[i], :] -- the "recoloured" points of the colour-quantisation use case
__global__ void __launch_bounds__(256) gather_centroids_kernel(const float* __restrict__ cent, const int32_t* __restrict__ labels,
                                                               int64_t n, int D, int K, float* __restrict__ out) {
    const int64_t total = n * D;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int64_t i = e / D;
        const int j = (int)(e - i * D);
        const int l = labels[i];
        out[e] = ((unsigned)l < (unsigned)K) ? cent[(int64_t)l * D + j] : __int_as_float(0x7fc00000);
    }
}

int launch_gather_centroids(kmc_ctx* c, const int32_t* labels, float* out, cudaStream_t s) {
    const int64_t total = c->n * c->D;
    if (total <= 0) return 0;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)c->sm_count * 16;
    if (blocks > cap) blocks = cap;
    gather_centroids_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->centroids(), labels, c->n, (int)c->D, (int)c->K, out);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

// max |x| over a flat fp32 array (NaNs ignored), folded into *out_bits with atomicMax on the
// bit pattern (non-negative floats order like unsigned ints).
__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ x, int64_t count,
                                                     unsigned int* __restrict__ out_bits) {
    float m = 0.f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    const int64_t nvec = vec_ok ? count / 4 : 0;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    for (int64_t i = tid; i < nvec; i += stride) {
        float4 v = ld_stream_f4(x4 + i);
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
    }
    for (int64_t i = nvec * 4 + tid; i < count; i += stride) m = fmaxf(m, fabsf(x[i]));
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.f && m < INFINITY) atomicMax(out_bits, __float_as_uint(m));
}

int launch_absmax(kmc_ctx* c, const float* x, int64_t count, unsigned int* out_bits, cudaStream_t s) {
    if (count <= 0) return 0;
    int64_t blocks = (count / 4 + 255) / 256 + 1;
    const int64_t cap = (int64_t)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    absmax_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, count, out_bits);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

// One warp per row: X'_hi = fp16(x * sigma), X'_lo = fp16(x * sigma - X'_hi) (both at the same
// scale), zero padding up to D_pad, and xnorm'[row] = |x * sigma| rounded up.
// Rows [row0, n).  keep_scale != 0: use the scale already in stats->sigma_x (rows arriving in pieces must all
// be split at the scale chosen for the first piece).
__global__ void __launch_bounds__(256) split_points_kernel(const float* __restrict__ x, int64_t row0, int64_t n, int D,
                                                           int D_pad, __half* __restrict__ xhi,
                                                           __half* __restrict__ xlo,
                                                           float* __restrict__ xnorm, __half* __restrict__ xext,
                                                           DeviceStats* stats, int keep_scale) {
    const float sigma = keep_scale ? stats->sigma_x : scale_from_absmax(__uint_as_float(stats->x_absmax_bits));
    if (!keep_scale && blockIdx.x == 0 && threadIdx.x == 0) stats->sigma_x = sigma;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = row0 + warp; row < n; row += nwarps) {
        const float* xr = x + row * (int64_t)D;
        __half* hr = xhi + row * (int64_t)D_pad;
        __half* lr = xlo + row * (int64_t)D_pad;
        float ss = 0.f, ss_lo = 0.f;
        for (int j = lane; j < D_pad; j += 32) {
            float v = (j < D) ? xr[j] * sigma : 0.f;
            __half h = __float2half_rn(v);
            float r = v - __half2float(h);          // exact
            hr[j] = h;
            lr[j] = __float2half_rn(r);
            ss = fmaf(v, v, ss);
            ss_lo = fmaf(r, r, ss_lo);
        }
        ss = warp_sum(ss);
        ss_lo = warp_sum(ss_lo);
        if (lane == 0) {
            xnorm[2 * row] = sqrtf(ss) * 1.001f;
            xnorm[2 * row + 1] = sqrtf(ss_lo) * 1.001f;
            // fp16-accumulator tier: the row's 16-column extension of the A operand for the bias MMA step --
            // [a_0..a_5 | r_0..r_5 | 0 0 0 0], a = kBiasConst (they meet the fp16 pieces of S|c_k|^2 in the centroid's
            // bias row) and sum_i a_i r_i = |X'_row|^2 (it meets kappa * a_i there: the MMA adds S|x|^2 to every score
            // of the row, so the scores of near centroids are small numbers with fine fp16 spacing)
            __align__(16) __half e[16];
            double r = (double)ss;
#pragma unroll
            for (int i = 0; i < kBiasPieces; ++i) {
                e[i] = __float2half_rn(kBiasConst[i]);
                float q = (float)(r / (double)kBiasConst[i]);
                q = fminf(fmaxf(q, -65504.f), 65504.f);
                const __half h = __float2half_rn(q);
                e[kBiasPieces + i] = h;
                r -= (double)kBiasConst[i] * (double)__half2float(h);
            }
#pragma unroll
            for (int i = 2 * kBiasPieces; i < 16; ++i) e[i] = __float2half_rn(0.f);
            uint4* dst = reinterpret_cast<uint4*>(xext + row * 16);
            dst[0] = *reinterpret_cast<const uint4*>(e);
            dst[1] = *reinterpret_cast<const uint4*>(e + 8);
        }
    }
}

int launch_split_points(kmc_ctx* c, cudaStream_t s) { return launch_split_points_range(c, 0, c->n, false, s); }

int launch_split_points_range(kmc_ctx* c, int64_t r0, int64_t r1, bool keep_scale, cudaStream_t s) {
    int64_t blocks = (r1 - r0 + 7) / 8;
    const int64_t cap = (int64_t)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    split_points_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->points, r0, r1, (int)c->D, (int)c->D_pad,
                                                         c->at<__half>(c->off_xhi), c->at<__half>(c->off_xlo),
                                                         c->at<float>(c->off_xnorm), c->at<__half>(c->off_xext), c->stats(),
                                                         keep_scale ? 1 : 0);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

}  // namespace kmc
