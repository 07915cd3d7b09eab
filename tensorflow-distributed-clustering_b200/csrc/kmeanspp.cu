// Seeded k-means++ initialisation on the device (D^2 sampling).
//
// The reference draws its initial centres with sklearn's private, unseeded
// `k_means_._init_centroids(data_batch, K, init='k-means++')` on the host
// (scripts/distribuitedClustering.py:191).  This is the plain k-means++ of Arthur & Vassilvitskii on
// the GPU: centre 0 is a uniformly drawn row; centre t is row i with probability
// mind[i] / sum(mind), mind[i] = min over the chosen centres of |x_i - c|^2 (exact fp32 difference
// form).  Per round one streaming pass updates mind and leaves an fp64 partial sum per block of 1024
// rows; a single-CTA kernel then locates the sampled row with two fp64 scans (over the block sums, then
// inside the block) and copies it out.  The K uniform numbers come from a splitmix64 stream of the
// seed generated on the host, so a run is reproducible and can be restated on the CPU
// (oracle/kmeanspp_oracle.py).
#include "kmc_common.cuh"
#include "kmc_device.cuh"

#include <vector>

namespace kmc {

constexpr int kPPBlockRows = 1024;

// mind[i] = min(mind[i], |x_i - c|^2) (first round: = |x_i - c|^2), block_sum[b] = sum of the block's mind
__global__ void __launch_bounds__(256) pp_update_kernel(const float* __restrict__ x, int64_t n, int D,
                                                        const float* __restrict__ centre, float* __restrict__ mind,
                                                        double* __restrict__ block_sum, int first) {
    extern __shared__ float cs[];
    __shared__ double wsum[8];
    for (int j = threadIdx.x; j < D; j += blockDim.x) cs[j] = centre[j];
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.x * kPPBlockRows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double acc = 0.0;
    // one warp per row, lanes over the columns
    for (int r = warp; r < kPPBlockRows; r += 8) {
        const int64_t row = r0 + r;
        if (row >= n) break;
        const float* xr = x + row * D;
        float d = 0.f;
        for (int j = lane; j < D; j += 32) {
            const float t = xr[j] - cs[j];
            d = fmaf(t, t, d);
        }
        d = warp_sum(d);
        if (lane == 0) {
            const float m = first ? d : fminf(mind[row], d);
            mind[row] = m;
            acc += (double)m;
        }
    }
    if (lane == 0) wsum[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += wsum[w];
        block_sum[blockIdx.x] = s;
    }
}

// inclusive scan of one double per thread over a 1024-thread CTA
__device__ double block_scan_1024(double v, double* sh) {
    const int t = threadIdx.x;
    sh[t] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const double a = (t >= o) ? sh[t - o] : 0.0;
        __syncthreads();
        sh[t] += a;
        __syncthreads();
    }
    return sh[t];
}

// pick the row whose cumulative mind first reaches u * total and copy it to `out`
__global__ void __launch_bounds__(1024) pp_select_kernel(const float* __restrict__ x, int64_t n, int D,
                                                         const float* __restrict__ mind,
                                                         const double* __restrict__ block_sum, int64_t n_blocks,
                                                         const double* __restrict__ uniforms, int t,
                                                         float* __restrict__ out, int64_t* __restrict__ chosen) {
    __shared__ double sh[1024];
    __shared__ double s_before;
    __shared__ long long s_block, s_row;
    const int tid = threadIdx.x;
    // thread i owns the contiguous slice [i*per, (i+1)*per) of the block sums
    const int64_t per = (n_blocks + 1023) / 1024;
    const int64_t b0 = (int64_t)tid * per, b1 = min(n_blocks, b0 + per);
    double mine = 0.0;
    for (int64_t b = b0; b < b1; ++b) mine += block_sum[b];
    const double incl = block_scan_1024(mine, sh);
    const double total = sh[1023];
    const double target = uniforms[t] * total;
    if (tid == 0) { s_block = -1; s_row = -1; }
    __syncthreads();
    // first slice whose inclusive sum reaches the target (total == 0: all rows coincide with a centre)
    const double excl = incl - mine;
    if (total > 0.0 && incl >= target && excl < target) {
        double run = excl;
        long long pick = b1 - 1;
        for (int64_t b = b0; b < b1; ++b) {
            if (run + block_sum[b] >= target) { pick = b; break; }
            run += block_sum[b];
        }
        s_block = pick;
        s_before = run;
    }
    __syncthreads();
    // degenerate (all distances zero) or rounding at the very end.  The flag is read into a register before
    // anyone rewrites it and the barriers below are taken by the whole CTA.
    const bool degenerate = s_block < 0;
    __syncthreads();
    if (degenerate && tid == 0) {
        s_block = (total > 0.0) ? n_blocks - 1 : 0;
        s_before = (total > 0.0) ? total - block_sum[n_blocks - 1] : 0.0;
    }
    __syncthreads();
    const int64_t base = (int64_t)s_block * kPPBlockRows;
    const int64_t row = base + tid;
    const double m = (row < n) ? (double)mind[row] : 0.0;
    const double incl_r = block_scan_1024(m, sh);
    const double want = target - s_before;
    if (m > 0.0 && incl_r >= want && incl_r - m < want) s_row = row;
    __syncthreads();
    if (s_row < 0 && tid == 0) {             // fall back to the last row of the block with a positive weight
        long long pick = base;
        for (int i = 0; i < kPPBlockRows && base + i < n; ++i)
            if (mind[base + i] > 0.f) pick = base + i;
        s_row = pick;
    }
    __syncthreads();
    for (int j = tid; j < D; j += 1024) out[j] = x[(int64_t)s_row * D + j];
    if (tid == 0) chosen[t] = s_row;
}

static uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

}  // namespace kmc

using namespace kmc;

extern "C" int64_t kmc_kmeanspp_scratch_bytes(int64_t n_rows, int64_t K) {
    if (n_rows < 0 || K < 1) return 0;
    const int64_t n_blocks = (n_rows + kPPBlockRows - 1) / kPPBlockRows;
    auto up = [](int64_t v) { return (v + 255) & ~(int64_t)255; };
    return up(n_rows * 4) + up(n_blocks * 8) + up(K * 8) + up(K * 8);
}

extern "C" int kmc_kmeanspp(int device, const float* dev_points, int64_t n_rows, int64_t D, int64_t K,
                            uint64_t seed, float* dev_centroids, int64_t* host_chosen_rows, void* dev_scratch,
                            int64_t scratch_bytes, void* stream) {
    if (!dev_points || !dev_centroids || !dev_scratch) KMC_FAIL(KMC_E_INVALID, "null argument");
    if (n_rows < 1 || D < 1 || K < 1 || K > n_rows) KMC_FAIL(KMC_E_INVALID, "need 1 <= K <= n_rows and D >= 1");
    if (scratch_bytes < kmc_kmeanspp_scratch_bytes(n_rows, K))
        KMC_FAIL(KMC_E_NOMEM, "scratch of %lld bytes given, %lld needed", (long long)scratch_bytes,
                 (long long)kmc_kmeanspp_scratch_bytes(n_rows, K));
    KMC_CUDA(cudaSetDevice(device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int64_t n_blocks = (n_rows + kPPBlockRows - 1) / kPPBlockRows;
    auto up = [](int64_t v) { return (v + 255) & ~(int64_t)255; };
    char* base = static_cast<char*>(dev_scratch);
    float* mind = reinterpret_cast<float*>(base);
    double* block_sum = reinterpret_cast<double*>(base + up(n_rows * 4));
    double* uniforms = reinterpret_cast<double*>(base + up(n_rows * 4) + up(n_blocks * 8));
    int64_t* chosen = reinterpret_cast<int64_t*>(base + up(n_rows * 4) + up(n_blocks * 8) + up(K * 8));

    std::vector<double> u((size_t)K);
    uint64_t st = seed;
    for (int64_t t = 0; t < K; ++t) u[(size_t)t] = (double)(splitmix64(st) >> 11) * (1.0 / 9007199254740992.0);
    KMC_CUDA(cudaMemcpyAsync(uniforms, u.data(), (size_t)K * 8, cudaMemcpyHostToDevice, s));
    KMC_CUDA(cudaStreamSynchronize(s));      // u lives on this stack frame

    // centre 0: uniform row
    int64_t first = (int64_t)(u[0] * (double)n_rows);
    if (first >= n_rows) first = n_rows - 1;
    KMC_CUDA(cudaMemcpyAsync(dev_centroids, dev_points + first * D, (size_t)D * 4, cudaMemcpyDeviceToDevice, s));
    KMC_CUDA(cudaMemcpyAsync(chosen, &first, 8, cudaMemcpyHostToDevice, s));
    KMC_CUDA(cudaStreamSynchronize(s));
    const size_t sm = (size_t)D * 4;
    if (sm > 48 * 1024) KMC_FAIL(KMC_E_UNSUPPORTED, "k-means++ supports D <= 12288");
    for (int64_t t = 1; t < K; ++t) {
        pp_update_kernel<<<(unsigned)n_blocks, 256, sm, s>>>(dev_points, n_rows, (int)D, dev_centroids + (t - 1) * D, mind,
                                                            block_sum, t == 1 ? 1 : 0);
        pp_select_kernel<<<1, 1024, 0, s>>>(dev_points, n_rows, (int)D, mind, block_sum, n_blocks, uniforms, (int)t,
                                            dev_centroids + t * D, chosen);
    }
    KMC_CUDA(cudaGetLastError());
    if (host_chosen_rows) {
        KMC_CUDA(cudaMemcpyAsync(host_chosen_rows, chosen, (size_t)K * 8, cudaMemcpyDeviceToHost, s));
        KMC_CUDA(cudaStreamSynchronize(s));
    }
    return 0;
}
