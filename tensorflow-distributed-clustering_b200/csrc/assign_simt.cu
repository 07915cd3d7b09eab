// Exact fp32 assignment for any (K, D): register-tiled difference-form distances
// d[i,k] = sum_j (x_ij - c_kj)^2 and a lowest-index argmin.  Two uses:
//   * the whole shard, for shapes neither the small path nor the tensor path covers;
//   * the refine pass of the tensor path: only the rows whose two best approximate scores are
//     closer than the tensor kernel's error bound (a device-side worklist).
//
// Replaces reference scripts/distribuitedClustering.py:221-234 (the [N,K,M] subtract / square /
// reduce_sum / argmin sub-graph) without forming the distance matrix.
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

constexpr int kSimtBM = 64, kSimtBN = 64, kSimtBK = 16, kSimtThreads = 256, kSimtPad = 4;

__global__ void __launch_bounds__(kSimtThreads)
simt_assign_kernel(const float* __restrict__ x, int64_t n_rows_total, int D, const float* __restrict__ cent,
                   int K, const int32_t* __restrict__ worklist, const unsigned long long* __restrict__ n_work_dev,
                   int64_t n_work_host, int32_t* __restrict__ labels, unsigned long long* __restrict__ over_stat) {
    __shared__ __align__(16) float Xs[kSimtBK][kSimtBM + kSimtPad];
    __shared__ __align__(16) float Cs[kSimtBK][kSimtBN + kSimtPad];
    __shared__ int64_t rowidx[kSimtBM];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    int64_t n_work = n_work_dev ? (int64_t)(*n_work_dev) : n_work_host;
    if (n_work > n_rows_total) n_work = n_rows_total;
    if (over_stat && blockIdx.x == 0 && tid == 0 && n_work > 0) atomicAdd(over_stat, (unsigned long long)n_work);
    const int64_t n_tiles = (n_work + kSimtBM - 1) / kSimtBM;
    const int lrow = tid >> 2;            // 0..63: row (or centroid) this thread stages
    const int lcol = (tid & 3) * 4;       // 0,4,8,12: first of 4 staged columns

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        __syncthreads();
        if (tid < kSimtBM) {
            const int64_t w = tile * kSimtBM + tid;
            rowidx[tid] = (w < n_work) ? (worklist ? (int64_t)worklist[w] : w) : -1;
        }
        __syncthreads();
        const int64_t my_row = rowidx[lrow];
        float best[4];
        int bk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { best[i] = INFINITY; bk[i] = 0x7fffffff; }

        for (int n0 = 0; n0 < K; n0 += kSimtBN) {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
            for (int k0 = 0; k0 < D; k0 += kSimtBK) {
                __syncthreads();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int col = k0 + lcol + q;
                    float xv = 0.f, cv = 0.f;
                    if (col < D) {
                        if (my_row >= 0) xv = x[my_row * (int64_t)D + col];
                        const int kc = n0 + lrow;
                        if (kc < K) cv = cent[(int64_t)kc * D + col];
                    }
                    Xs[lcol + q][lrow] = xv;
                    Cs[lcol + q][lrow] = cv;
                }
                __syncthreads();
#pragma unroll
                for (int kk = 0; kk < kSimtBK; ++kk) {
                    const float4 a = *reinterpret_cast<const float4*>(&Xs[kk][ty * 4]);
                    const float4 b = *reinterpret_cast<const float4*>(&Cs[kk][tx * 4]);
                    const float av[4] = {a.x, a.y, a.z, a.w};
                    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float t = av[i] - bv[j];
                            acc[i][j] = fmaf(t, t, acc[i][j]);
                        }
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = n0 + tx * 4 + j;
                if (k < K) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (acc[i][j] < best[i]) { best[i] = acc[i][j]; bk[i] = k; }
                }
            }
        }
        // lexicographic (distance, index) min across the 16 threads that share a row
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float d = best[i];
            int k = bk[i];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float od = __shfl_xor_sync(0xffffffffu, d, o);
                const int ok = __shfl_xor_sync(0xffffffffu, k, o);
                if (od < d || (od == d && ok < k)) { d = od; k = ok; }
            }
            if (tx == 0) {
                const int64_t r = rowidx[ty * 4 + i];
                if (r >= 0) labels[r] = (k == 0x7fffffff) ? 0 : k;
            }
        }
    }
}

int launch_simt_assign(kmc_ctx* c, const int32_t* worklist, const unsigned long long* n_work_dev,
                       int64_t n_work_host, int32_t* labels_out, cudaStream_t s, unsigned long long* over_stat) {
    int64_t blocks;
    if (n_work_dev) {
        blocks = (int64_t)c->sm_count * 4;
    } else {
        blocks = (n_work_host + kSimtBM - 1) / kSimtBM;
        const int64_t cap = (int64_t)c->sm_count * 8;
        if (blocks > cap) blocks = cap;
    }
    if (blocks < 1) return 0;
    simt_assign_kernel<<<(unsigned)blocks, kSimtThreads, 0, s>>>(c->points, c->n, (int)c->D, c->centroids(),
                                                                 (int)c->K, worklist, n_work_dev, n_work_host,
                                                                 labels_out, over_stat);
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

}  // namespace kmc
