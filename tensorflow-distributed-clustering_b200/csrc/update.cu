// Centroid update + convergence statistics, and (tensor path) the per-iteration preparation of
// the split fp16 centroid planes.
//
// Replaces reference scripts/distribuitedClustering.py:257-263: add_n of the per-GPU partials
// (done before this kernel by the host's all-reduce), tile/reshape/div/transpose and the assign
// of `global_centroids` on /cpu:0.  The division is done in fp64 like the reference and rounded
// once to fp32.  A globally empty cluster keeps its centroid (KMC_EMPTY_KEEP) or becomes NaN
// (KMC_EMPTY_NAN, the reference's 0/0).
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

__global__ void __launch_bounds__(256) update_kernel(float* __restrict__ cent, float* __restrict__ cent_prev,
                                                     const double* __restrict__ partials,
                                                     int K, int D, int empty_nan, DeviceStats* stats,
                                                     double* __restrict__ shift_slots) {
    const int64_t total = (int64_t)K * D;
    double shift = 0.0;
    float amax = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i / D);
        const double cnt = partials[total + k];
        const float old = cent[i];
        cent_prev[i] = old;              // the centroids the last pass was assigned against (kmc_last_labels)
        float nw;
        if (cnt > 0.0) nw = (float)(partials[i] / cnt);
        else nw = empty_nan ? __int_as_float(0x7fc00000) : old;
        cent[i] = nw;
        const double d = (double)nw - (double)old;
        if (d == d) shift += d * d;
        amax = fmaxf(amax, fabsf(nw));
    }
    // shift2 in a fixed summation order (lanes by butterfly, warps and CTAs in index order, the last CTA to finish
    // adds the per-CTA slots): every GPU computes bit-identical values from identical inputs, so ranks that test
    // shift2 <= tol on their own take the same decision
    __shared__ double wsum[8];
    shift = warp_sum(shift);
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0) {
        wsum[threadIdx.x >> 5] = shift;
        if (amax > 0.f && amax < INFINITY) atomicMax(&stats->c_absmax_bits, __float_as_uint(amax));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += wsum[w];
        shift_slots[blockIdx.x] = s;
        __threadfence();
        if (atomicAdd(&stats->upd_done, 1u) == gridDim.x - 1) {
            __threadfence();
            double total = 0.0;
            for (unsigned b = 0; b < gridDim.x; ++b) total += *(volatile double*)(shift_slots + b);
            stats->shift2 = total;
            stats->upd_done = 0u;
        }
    }
}

int launch_update(kmc_ctx* c, cudaStream_t s) {
    const int64_t total = c->K * c->D;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = kUpdateMaxBlocks < (int64_t)c->sm_count * 4 ? kUpdateMaxBlocks : (int64_t)c->sm_count * 4;
    blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
    // c_absmax restarts at zero for every update (shift2 is overwritten by the kernel)
    KMC_CUDA(cudaMemsetAsync(&c->stats()->c_absmax_bits, 0, sizeof(unsigned int), s));
    update_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->centroids(), c->centroids_prev(), c->partials(), (int)c->K, (int)c->D,
                                                   (c->flags & KMC_EMPTY_NAN) ? 1 : 0, c->stats(),
                                                   c->at<double>(c->off_shift_slots));
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

__global__ void __launch_bounds__(256) centroid_absmax_kernel(const float* __restrict__ cent, int64_t total,
                                                              DeviceStats* stats) {
    float amax = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x)
        amax = fmaxf(amax, fabsf(cent[i]));
    amax = warp_max(amax);
    if ((threadIdx.x & 31) == 0 && amax > 0.f && amax < INFINITY)
        atomicMax(&stats->c_absmax_bits, __float_as_uint(amax));
}

int launch_centroid_absmax(kmc_ctx* c, cudaStream_t s) {
    const int64_t total = c->K * c->D;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)c->sm_count * 4;
    blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
    KMC_CUDA(cudaMemsetAsync(&c->stats()->c_absmax_bits, 0, sizeof(unsigned int), s));
    centroid_absmax_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->centroids(), total, c->stats());
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

// One warp per (padded) centroid row: C'_hi / C'_lo = split fp16 of (-2 * sigma_c * c), and
// cn'[k] = sigma_x * sigma_c * |c_k|^2 (3e38 for padding rows and NaN centroids, so they never
// win), plus max_k |C'_k| for the refine threshold.
// fp16-accumulator tier: a second, smaller centroid scale sigma_c16 = rho16 * sigma_c (power of two) such that no
// partial sum of a score can leave the fp16 range: S16 |x_i - c_k|^2 and every part of it is at most
// S16 * D * (xabs + cabs)^2 < 24000 with S16 = sigma_x * sigma_c16 (|v| <= sqrt(D) * absmax for rows and centroids).
__device__ __forceinline__ float rho16_for(float sigma_x, float sigma_c, float xabs, float cabs, int D) {
    const float bound = sigma_x * sigma_c * (float)D * (xabs + cabs) * (xabs + cabs);
    if (!(bound > 0.f) || !(bound < 3.0e38f)) return 1.f;
    int e;
    frexpf(bound / 24000.f, &e);          // bound / 24000 = f * 2^e, f in [0.5, 1)  ->  rho = 2^-e keeps bound * rho < 24000
    e = e < 0 ? 0 : (e > 60 ? 60 : e);
    return ldexpf(1.f, -e);
}

__global__ void __launch_bounds__(256) prep_centroids_kernel(const float* __restrict__ cent, int K, int D,
                                                             int K_pad, int D_pad, __half* __restrict__ chi,
                                                             __half* __restrict__ clo, float* __restrict__ cnorm,
                                                             __half* __restrict__ cbias, __half* __restrict__ c16,
                                                             __half* __restrict__ cbias16, DeviceStats* stats) {
    // 2*|c| < 2^14 after scaling: absmax of (2c)
    const float sigma_c = scale_from_absmax(2.f * __uint_as_float(stats->c_absmax_bits));
    const float sigma_x = stats->sigma_x;
    const float rho16 = rho16_for(sigma_x, sigma_c, __uint_as_float(stats->x_absmax_bits), __uint_as_float(stats->c_absmax_bits), D);
    if (blockIdx.x == 0 && threadIdx.x == 0) { stats->sigma_c = sigma_c; stats->rho16 = rho16; }
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    float nmax = 0.f, lomax = 0.f;
    for (int k = warp; k < K_pad; k += nwarps) {
        __half* hr = chi + (int64_t)k * D_pad;
        __half* lr = clo + (int64_t)k * D_pad;
        double ss = 0.0;
        float ss_scaled = 0.f;
        if (k < K) {
            for (int j = lane; j < D; j += 32) {
                const float cval = cent[(int64_t)k * D + j];
                const float v = -2.f * sigma_c * cval;
                ss += (double)cval * (double)cval;
                ss_scaled = fmaf(v, v, ss_scaled);
            }
        }
        ss = warp_sum(ss);
        ss_scaled = warp_sum(ss_scaled);
        float cn = (float)(ss * (double)sigma_x * (double)sigma_c);
        // padding rows, NaN / infinite centroids: zero planes and a huge (finite, so that the epilogue can
        // pack an index into its low bits) bias -> never chosen
        const bool dead = (k >= K) || !(cn < INFINITY) || !(ss_scaled < INFINITY);
        float ss_lo = 0.f;
        for (int j = lane; j < D_pad; j += 32) {
            float v = 0.f;
            if (!dead && j < D) v = -2.f * sigma_c * cent[(int64_t)k * D + j];
            const __half h = __float2half_rn(v);
            const float r = v - __half2float(h);
            hr[j] = h;
            lr[j] = __float2half_rn(r);
            ss_lo = fmaf(r, r, ss_lo);
            c16[(int64_t)k * D_pad + j] = __float2half_rn(v * rho16);      // = rho16 * h unless it goes subnormal
        }
        ss_lo = warp_sum(ss_lo);
        // bias plane: row k = fp16 pieces p_i with sum_i a_i * p_i = cn (a = kBiasConst), zero padded to one
        // 64-column chunk; the tensor kernel multiplies it with a constant tile holding the a_i, so the MMA
        // itself adds S*|c_k|^2 to every score.  Dead rows saturate (8.6e9, above every real score).
        {
            __half* br = cbias + (int64_t)k * kBiasCols;
            if (lane < kBiasCols) br[lane] = __float2half_rn(0.f);
            __syncwarp();
            if (lane == 0) {
                double r = (double)cn;
                for (int i = 0; i < kBiasPieces; ++i) {
                    float q = dead ? ((i < 4) ? 65504.f : 0.f) : (float)(r / (double)kBiasConst[i]);
                    q = fminf(fmaxf(q, -65504.f), 65504.f);
                    const __half h = __float2half_rn(q);
                    br[i] = h;
                    r -= (double)kBiasConst[i] * (double)__half2float(h);
                }
                if (!dead) atomicMax(&stats->cn_max_bits, __float_as_uint(cn));
                // the same pieces at the scale of the fp16-accumulator tier; a dead row gets 49152 there, above every
                // real score of that tier (< 24000) and inside the fp16 range
                __half* br16 = cbias16 + (int64_t)k * kBiasCols;
                double r16 = dead ? 49152.0 : (double)cn * (double)rho16;
                for (int i = 0; i < kBiasPieces; ++i) {
                    float q = (float)(r16 / (double)kBiasConst[i]);
                    q = fminf(fmaxf(q, -65504.f), 65504.f);
                    const __half h = __float2half_rn(q);
                    br16[i] = h;
                    r16 -= (double)kBiasConst[i] * (double)__half2float(h);
                }
            }
            __syncwarp();
            // columns 6..11: kappa * a_i, the partners of the row's pieces of |X'|^2 (kappa = S16 / sigma_x^2)
            if (lane >= kBiasPieces && lane < kBiasCols) {
                float q = 0.f;
                if (lane < 2 * kBiasPieces) q = fminf(kBiasConst[lane - kBiasPieces] * (rho16 * sigma_c / sigma_x), 65504.f);
                cbias16[(int64_t)k * kBiasCols + lane] = __float2half_rn(q);
            }
        }
        if (lane == 0) {
            cnorm[k] = dead ? 3.0e38f : cn;
            if (!dead) {
                nmax = fmaxf(nmax, sqrtf(ss_scaled) * 1.001f);
                lomax = fmaxf(lomax, sqrtf(ss_lo) * 1.001f);
            }
        }
    }
    if (lane == 0 && nmax > 0.f && nmax < INFINITY) atomicMax(&stats->c_normmax_bits, __float_as_uint(nmax));
    if (lane == 0 && lomax > 0.f && lomax < INFINITY) atomicMax(&stats->c_lonormmax_bits, __float_as_uint(lomax));
}

int launch_prep_centroids(kmc_ctx* c, cudaStream_t s) {
    int64_t blocks = (c->K_pad + 7) / 8;
    const int64_t cap = (int64_t)c->sm_count * 4;
    blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
    KMC_CUDA(cudaMemsetAsync(&c->stats()->c_normmax_bits, 0, sizeof(unsigned int), s));
    KMC_CUDA(cudaMemsetAsync(&c->stats()->c_lonormmax_bits, 0, sizeof(unsigned int), s));
    KMC_CUDA(cudaMemsetAsync(&c->stats()->cn_max_bits, 0, sizeof(unsigned int), s));
    __half* chi = c->at<__half>(c->off_cplanes);
    __half* clo = chi + c->K_pad * c->D_pad;
    prep_centroids_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->centroids(), (int)c->K, (int)c->D, (int)c->K_pad,
                                                           (int)c->D_pad, chi, clo, c->at<float>(c->off_cnorm),
                                                           c->at<__half>(c->off_cbias), c->at<__half>(c->off_cplane16),
                                                           c->at<__half>(c->off_cbias16), c->stats());
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

}  // namespace kmc
