// Exact re-decision of the rows the tensor kernel could not certify (its two best approximate
// scores are closer than its error bound).  One warp per row; candidates are pruned with the
// triangle inequality before any exact distance is computed:
//     |c_k - c_a| >= 2 |x - c_a|   ==>   |x - c_k| >= |x - c_a|      (a = the tensor kernel's pick)
// so only centroids with ccdist[a][k] < 4 d(x, c_a) can win.  ccdist is the K x K matrix of squared
// centroid distances, rebuilt once per iteration (K*K*D FMAs, negligible next to N*K*D).  The
// surviving candidates get the reference's exact fp32 difference-form distance and a
// lowest-index argmin, so the outcome equals a full exact scan.
//
// Part of the replacement of reference scripts/distribuitedClustering.py:221-234.
#include "kmc_common.cuh"
#include "kmc_device.cuh"

namespace kmc {

// ccdist[i][j] = sum_d (c_i[d] - c_j[d])^2, fp32, 16x16 output tile per CTA
__global__ void __launch_bounds__(256) ccdist_kernel(const float* __restrict__ cent, int K, int D,
                                                     float* __restrict__ cc) {
    __shared__ float A[16][17], B[16][17];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int i = blockIdx.y * 16 + ty, j = blockIdx.x * 16 + tx;
    float acc = 0.f;
    for (int d0 = 0; d0 < D; d0 += 16) {
        const int ia = blockIdx.y * 16 + ty, ib = blockIdx.x * 16 + ty;
        A[ty][tx] = (ia < K && d0 + tx < D) ? cent[(int64_t)ia * D + d0 + tx] : 0.f;
        B[ty][tx] = (ib < K && d0 + tx < D) ? cent[(int64_t)ib * D + d0 + tx] : 0.f;
        __syncthreads();
#pragma unroll
        for (int d = 0; d < 16; ++d) {
            const float t = A[ty][d] - B[tx][d];
            acc = fmaf(t, t, acc);
        }
        __syncthreads();
    }
    if (i < K && j < K) cc[(int64_t)i * K + j] = acc;
}

__global__ void nbrsort_kernel(float* __restrict__ cc, int32_t* __restrict__ nbr, int K);

int launch_ccdist(kmc_ctx* c, cudaStream_t s) {
    const int K = (int)c->K;
    dim3 grid((K + 15) / 16, (K + 15) / 16);
    ccdist_kernel<<<grid, 256, 0, s>>>(c->centroids(), K, (int)c->D, c->at<float>(c->off_ccdist));
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    if (c->off_nbr >= 0) {
        int P = 1;
        while (P < K) P <<= 1;
        const size_t sm = (size_t)P * sizeof(unsigned long long);
        if (sm > 48 * 1024)
            KMC_CUDA(cudaFuncSetAttribute(nbrsort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        nbrsort_kernel<<<K, 512, sm, s>>>(c->at<float>(c->off_ccdist), c->at<int32_t>(c->off_nbr), K);
        KMC_CUDA(cudaGetLastError());
        c->launches++;
    }
    return 0;
}

// Sort every row of ccdist ascending (one CTA per row, bitonic sort of (distance, index) keys in
// shared memory), in place, with the permutation in nbr: the refine kernel then walks the
// neighbours of its anchor nearest-first and stops at the pruning radius, touching a few entries
// instead of the whole row.  K <= 8192 (64 KB of keys).

__global__ void __launch_bounds__(512) nbrsort_kernel(float* __restrict__ cc, int32_t* __restrict__ nbr, int K) {
    extern __shared__ unsigned long long keys[];
    const int a = blockIdx.x;
    int P = 1;
    while (P < K) P <<= 1;
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        unsigned long long key = ~0ull;
        if (i < K) {
            const float v = cc[(int64_t)a * K + i];
            const unsigned int b = (v == v) ? __float_as_uint(fmaxf(v, 0.f)) : 0x7fffffffu;   // NaN last
            key = ((unsigned long long)b << 32) | (unsigned int)i;
        }
        keys[i] = key;
    }
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
                const int lo = 2 * t - (t & (stride - 1));
                const int hi = lo + stride;
                const bool up = ((lo & size) == 0);
                const unsigned long long x0 = keys[lo], x1 = keys[hi];
                if ((x0 > x1) == up) { keys[lo] = x1; keys[hi] = x0; }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        const unsigned long long key = keys[i];
        const unsigned int b = (unsigned int)(key >> 32);
        cc[(int64_t)a * K + i] = (b == 0x7fffffffu) ? __int_as_float(0x7fc00000) : __uint_as_float(b);
        nbr[(int64_t)a * K + i] = (int32_t)(key & 0xffffffffu);
    }
}

template <int NPL, bool SORTED>   // floats of the row held per lane: D <= 32 * NPL
__global__ void __launch_bounds__(256)
refine_kernel(const float* __restrict__ x, int D, const float* __restrict__ cent, int K,
              const float* __restrict__ cc, const int32_t* __restrict__ nbr,
              const int32_t* __restrict__ worklist,
              const unsigned long long* __restrict__ n_work_dev, int32_t* __restrict__ labels) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_work = (int64_t)(*n_work_dev);
    for (int64_t w = warp; w < n_work; w += nwarps) {
        const int64_t row = worklist[w];
        const int32_t lab = labels[row];
        int a = anchor_first(lab);
        const int a_alt = anchor_second(lab);
        float xr[NPL];
#pragma unroll
        for (int q = 0; q < NPL; ++q) {
            const int d = lane + 32 * q;
            xr[q] = d < D ? x[row * D + d] : 0.f;
        }
        auto dist_to = [&](int k) {
            float acc = 0.f;
#pragma unroll
            for (int q = 0; q < NPL; ++q) {
                const int d = lane + 32 * q;
                const float t = xr[q] - (d < D ? cent[(int64_t)k * D + d] : 0.f);
                acc = fmaf(t, t, acc);
            }
            return warp_sum(acc);
        };
        if (a_alt >= 0 && a_alt < K && dist_to(a_alt) < dist_to(a)) a = a_alt;     // two anchor candidates: the nearer
        float best = dist_to(a);
        int bk = a;
        // a small relative margin absorbs the fp32 rounding of ccdist and of `best`
        const float lim = 4.f * best * 1.0001f + 1e-30f;
        const float* ccrow = cc + (int64_t)a * K;
        for (int k0 = 0; k0 < K; k0 += 32) {
            const int j = k0 + lane;
            const float ccv = (j < K) ? ccrow[j] : INFINITY;
            const int k = SORTED ? ((j < K) ? nbr[(int64_t)a * K + j] : -1) : j;
            const bool cand = (j < K) && (k != a) && (ccv < lim);
            unsigned m = __ballot_sync(0xffffffffu, cand);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const int kk = SORTED ? __shfl_sync(0xffffffffu, k, src) : (k0 + src);
                const float d = dist_to(kk);
                if (d < best || (d == best && kk < bk)) { best = d; bk = kk; }
            }
            // sorted rows: once one entry is outside the pruning radius, all later ones are
            if (SORTED && __ballot_sync(0xffffffffu, !(ccv < lim)) != 0u) break;
        }
        if (lane == 0) labels[row] = bk;
    }
}

// Faster variant for sorted neighbour lists and D % 4 == 0: the warp evaluates FOUR candidates at a time
// (8 lanes per candidate, each lane a strided set of float4 slices of the row), starting with the first
// four entries of the anchor's list (entry 0 is the anchor itself), so the usual row -- anchor plus one
// or two siblings inside the pruning radius -- costs one batch.
// Walk-length feedback: every refine launch records how many exact distances it evaluated per row; the next
// launch (both kernels are enqueued, one returns at once) runs the warp-per-row kernel when the previous walks
// were long (far-from-converged centroids: hundreds of candidates inside the pruning radius) and the
// 8-lanes-per-row kernel when they were short (converged blobs: two candidates).  Unknown (first launch) = long.
__device__ __forceinline__ bool long_walks_expected(const DeviceStats* stats, int prev_slot) {
    const unsigned long long r = stats->walk[prev_slot][1], s = stats->walk[prev_slot][0];
    return r == 0ull || s > 6ull * r;
}

// Candidates a refine walk evaluates before the row is handed to the tile kernel: K / 8 (between 32 and 256).  A walk
// costs ~1.5 us of dependent loads per 4 candidates, the register-tiled kernel evaluates all K candidates of a row
// at full FFMA rate; measured at K = 1024, D = 128 the two meet near 128 candidates.
static int walk_cap_for(int64_t K) { return K / 8 < 32 ? 32 : (K / 8 > 256 ? 256 : (int)(K / 8)); }

template <int NV>   // float4 slices per lane: D <= 32 * NV * 4 / 4 ... lane t owns float4 indices t, t+8, ...
__global__ void __launch_bounds__(256)
refine4_kernel(const float* __restrict__ x, int D, const float* __restrict__ cent, int K,
               const float* __restrict__ cc, const int32_t* __restrict__ nbr,
               const int32_t* __restrict__ worklist, const unsigned long long* __restrict__ n_work_dev,
               int32_t* __restrict__ labels, DeviceStats* __restrict__ stats, int slot, int select,
               int32_t* __restrict__ overflow, unsigned long long* __restrict__ n_over, int kWalkCap) {
    if (select && !long_walks_expected(stats, slot ^ 1)) return;      // refine8_kernel runs this launch instead
    const int lane = threadIdx.x & 31, g = lane >> 3, t = lane & 7;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_work = (int64_t)(*n_work_dev);
    const int nvec = D >> 2;
    unsigned long long walked = 0, rows_done = 0;
    for (int64_t w = warp; w < n_work; w += nwarps) {
        ++rows_done;
        const int64_t row = worklist[w];
        const int32_t lab = labels[row];
        int a = anchor_first(lab);
        const int a_alt = anchor_second(lab);
        float4 xr[NV];
        const float4* xp = reinterpret_cast<const float4*>(x + row * D);
#pragma unroll
        for (int q = 0; q < NV; ++q) {
            const int v = t + 8 * q;
            xr[q] = v < nvec ? xp[v] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (a_alt >= 0 && a_alt < K) {          // two anchor candidates (warp-uniform): lane groups 0 / 1 measure them
            const int kc = (g & 1) ? a_alt : a;
            const float4* cp = reinterpret_cast<const float4*>(cent + (int64_t)kc * D);
            float acc = 0.f;
#pragma unroll
            for (int q = 0; q < NV; ++q) {
                const int v = t + 8 * q;
                if (v < nvec) {
                    const float4 cv = cp[v];
                    const float d0 = xr[q].x - cv.x, d1 = xr[q].y - cv.y, d2 = xr[q].z - cv.z, d3 = xr[q].w - cv.w;
                    acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
                }
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            const float da = __shfl_sync(0xffffffffu, acc, 0), dalt = __shfl_sync(0xffffffffu, acc, 8);
            if (dalt < da) a = a_alt;
        }
        const float* ccrow = cc + (int64_t)a * K;
        const int32_t* nbrow = nbr + (int64_t)a * K;
        float best = INFINITY, lim = INFINITY;
        int bk = 0x7fffffff;
        bool complete = false;
        int j0 = 0;
        for (; j0 < K; j0 += 4) {
            const int j = j0 + g;
            const float ccv = (j < K) ? ccrow[j] : INFINITY;
            const int k = (j < K) ? nbrow[j] : -1;
            float acc = 0.f;
            if (k >= 0) {
                const float4* cp = reinterpret_cast<const float4*>(cent + (int64_t)k * D);
#pragma unroll
                for (int q = 0; q < NV; ++q) {
                    const int v = t + 8 * q;
                    if (v < nvec) {
                        const float4 cv = cp[v];
                        float d0 = xr[q].x - cv.x, d1 = xr[q].y - cv.y, d2 = xr[q].z - cv.z, d3 = xr[q].w - cv.w;
                        acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
                    }
                }
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            if (j0 == 0) {
                // entry 0 of the sorted list is the anchor (or an identical centroid): its distance sets the
                // pruning radius; a small relative margin absorbs the fp32 rounding of ccdist and of the distance
                const float d_anchor = __shfl_sync(0xffffffffu, acc, 0);
                lim = 4.f * d_anchor * 1.0001f + 1e-30f;
            }
            float d = (k >= 0 && (ccv < lim || j == 0)) ? acc : INFINITY;
            int kk = (d < INFINITY) ? k : 0x7fffffff;
            // lexicographic (distance, index) min over the four candidate groups
#pragma unroll
            for (int o = 8; o <= 16; o <<= 1) {
                const float od = __shfl_xor_sync(0xffffffffu, d, o);
                const int ok = __shfl_xor_sync(0xffffffffu, kk, o);
                if (od < d || (od == d && ok < kk)) { d = od; kk = ok; }
            }
            if (d < best || (d == best && kk < bk)) { best = d; bk = kk; }
            walked += 4;
            // sorted: once an entry of this batch is outside the radius, all later ones are
            if (__ballot_sync(0xffffffffu, !(ccv < lim)) != 0u) { complete = true; break; }
            // bounded walk: a row with more than kWalkCap candidates inside its pruning radius (data without
            // cluster structure, far-from-converged centroids) goes to the register-tiled exact kernel instead
            if (j0 + 4 >= kWalkCap && j0 + 4 < K) break;
        }
        complete = complete || j0 >= K;
        if (lane == 0) {
            if (complete) labels[row] = (bk == 0x7fffffff) ? a : bk;
            else overflow[atomicAdd(n_over, 1ull)] = (int32_t)row;
        }
    }
    if (lane == 0 && rows_done) {
        atomicAdd(&stats->walk[slot][0], walked);
        atomicAdd(&stats->walk[slot][1], rows_done);
    }
}

// Eight lanes per row, four rows per warp.  Most uncertified rows need two exact distances (the tensor
// kernel's pick and the centroid sharing its blob): entries 0 and 1 of the anchor's sorted list are evaluated
// together, then the walk continues four entries per step (independent accumulators, each predicated on the
// pruning radius) until the first entry outside the radius.  Same pruning rule, same lexicographic
// (distance, index) argmin and the same walk order as refine4_kernel, which spends a whole warp on one row.
template <int NV>
__device__ __forceinline__ float group_dist(const float4 (&xr)[NV], const float* __restrict__ cent, int k, int D,
                                            int nvec, int t) {
    float acc = 0.f;
    if (k >= 0) {
        const float4* cp = reinterpret_cast<const float4*>(cent + (int64_t)k * D);
#pragma unroll
        for (int q = 0; q < NV; ++q) {
            const int v = t + 8 * q;
            if (v < nvec) {
                const float4 cv = cp[v];
                const float d0 = xr[q].x - cv.x, d1 = xr[q].y - cv.y, d2 = xr[q].z - cv.z, d3 = xr[q].w - cv.w;
                acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
            }
        }
    }
    return acc;
}

__device__ __forceinline__ float group_sum8(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}

// Four CTAs per SM (64 registers): the kernel waits on the gathered point rows, and throughput follows the rows in
// flight -- 2.35 -> 1.82 ms for the 11 M refined rows of config 4 (ncu, one B200) against three CTAs at 70 registers.
// Tried on top of that and measured at no gain (1.75-2.0 ms each): resolving worklist entry -> label -> head of the
// neighbour list for 1024 rows at once into shared memory (two of the four dependent hops gone), evaluating the
// three nearest candidates together (2.3x fewer instructions), requesting the next row and label a step ahead
// (slower: spills).  What is left is the gather itself: 6.1 GB of 512-byte rows at 3.4 TB/s.
template <int NV>
__global__ void __launch_bounds__(256, 4)
refine8_kernel(const float* __restrict__ x, int D, const float* __restrict__ cent, int K,
               const float* __restrict__ cc, const int32_t* __restrict__ nbr,
               const int32_t* __restrict__ worklist, const unsigned long long* __restrict__ n_work_dev,
               int32_t* __restrict__ labels, DeviceStats* __restrict__ stats, int slot, int select,
               int32_t* __restrict__ overflow, unsigned long long* __restrict__ n_over, int kWalkCap) {
    if (select && long_walks_expected(stats, slot ^ 1)) return;       // refine4_kernel runs this launch instead
    const int lane = threadIdx.x & 31, g = lane >> 3, t = lane & 7;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_work = (int64_t)(*n_work_dev);
    const int nvec = D >> 2;
    unsigned long long walked = 0, rows_done = 0;
    for (int64_t w0 = warp * 4; w0 < n_work; w0 += nwarps * 4) {
        const int64_t w = w0 + g;
        const bool live = w < n_work;
        if (live) { ++rows_done; walked += 2; }
        const int64_t row = live ? worklist[w] : 0;
        const int32_t lab = live ? labels[row] : 0;
        int a = anchor_first(lab);
        int a_alt = anchor_second(lab);
        if (a_alt >= K) a_alt = -1;
        float4 xr[NV];
        const float4* xp = reinterpret_cast<const float4*>(x + row * D);
#pragma unroll
        for (int q = 0; q < NV; ++q) {
            const int v = t + 8 * q;
            xr[q] = (live && v < nvec) ? ld_stream_f4(xp + v) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (__any_sync(0xffffffffu, a_alt >= 0)) {      // a row with two anchor candidates: its lane group takes the nearer
            float da = group_dist<NV>(xr, cent, a_alt >= 0 ? a : -1, D, nvec, t);
            float dalt = group_dist<NV>(xr, cent, a_alt, D, nvec, t);
            da = group_sum8(da);
            dalt = group_sum8(dalt);
            if (a_alt >= 0 && dalt < da) a = a_alt;
        }
        const float* ccrow = cc + (int64_t)a * K;
        const int32_t* nbrow = nbr + (int64_t)a * K;
        float best = INFINITY, lim = INFINITY;
        int bk = 0x7fffffff;
        auto take = [&](float d, int k) {
            if (k >= 0 && (d < best || (d == best && k < bk))) { best = d; bk = k; }
        };
        // entries 0 (the anchor, or an identical centroid: its distance sets the pruning radius) and 1
        const int k0 = live ? nbrow[0] : -1;
        const int k1 = (live && K > 1) ? nbrow[1] : -1;
        const float c1 = (live && K > 1) ? ccrow[1] : INFINITY;
        const float c2 = (live && K > 2) ? ccrow[2] : INFINITY;
        float d0 = group_dist<NV>(xr, cent, k0, D, nvec, t);
        float d1 = group_dist<NV>(xr, cent, k1, D, nvec, t);
        d0 = group_sum8(d0);
        d1 = group_sum8(d1);
        if (live) lim = 4.f * d0 * 1.0001f + 1e-30f;
        take(d0, k0);
        if (c1 < lim) take(d1, k1);
        // sorted list: the first entry outside the radius ends the walk
        bool active = live && (c1 < lim) && (c2 < lim);
        bool over = false;
        for (int j = 2; __any_sync(0xffffffffu, active); j += 4) {
            int kk[4];
            bool inside = true;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int ju = j + u;
                inside = inside && active && ju < K && (ccrow[ju] < lim);
                kk[u] = inside ? nbrow[ju] : -1;
            }
            float dd[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) dd[u] = group_dist<NV>(xr, cent, kk[u], D, nvec, t);
#pragma unroll
            for (int u = 0; u < 4; ++u) dd[u] = group_sum8(dd[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u) take(dd[u], kk[u]);
            if (active) walked += 4;
            active = inside;
            if (active && j + 4 >= kWalkCap && j + 4 < K) { over = true; active = false; }     // bounded walk
        }
        if (live && t == 0) {
            if (over) overflow[atomicAdd(n_over, 1ull)] = (int32_t)row;
            else labels[row] = (bk == 0x7fffffff) ? a : bk;
        }
    }
    if (t == 0 && rows_done) {
        atomicAdd(&stats->walk[slot][0], walked);
        atomicAdd(&stats->walk[slot][1], rows_done);
    }
}

int launch_refine(kmc_ctx* c, int32_t* labels, const int32_t* worklist, const unsigned long long* nw,
                  int32_t* overflow, unsigned long long* n_over, cudaStream_t s) {
    const int blocks = c->sm_count * 8;
    const int npl = (int)((c->D + 31) / 32);
    const float* cc = c->at<float>(c->off_ccdist);
    const bool sorted = c->off_nbr >= 0;
    const int32_t* nbr = sorted ? c->at<int32_t>(c->off_nbr) : nullptr;
    if (sorted && c->D % 4 == 0 && c->D <= 256 && ((reinterpret_cast<uintptr_t>(c->points) & 15) == 0)) {
        const int nv = (int)((c->D / 4 + 7) / 8);
        // KMC_PARAM_REFINE_KERNEL = 4 | 8 pins one kernel; default: both enqueued, the device picks
        const int pin = c->refine_kernel;
        const int slot = (int)(c->refine_launches++ & 1u);
        const int cap = walk_cap_for(c->K);
        KMC_CUDA(cudaMemsetAsync(&c->stats()->walk[slot], 0, 4 * sizeof(unsigned long long), s));
#define KMC_REFINE4(N)                                                                                          \
    do {                                                                                                        \
        if (pin != 8)                                                                                           \
            refine4_kernel<N><<<blocks, 256, 0, s>>>(c->points, (int)c->D, c->centroids(), (int)c->K, cc, nbr,  \
                                                     worklist, nw, labels, c->stats(), slot, pin == 0, overflow, n_over, cap); \
        if (pin != 4)                                                                                           \
            refine8_kernel<N><<<blocks, 256, 0, s>>>(c->points, (int)c->D, c->centroids(), (int)c->K, cc, nbr,  \
                                                     worklist, nw, labels, c->stats(), slot, pin == 0, overflow, n_over, cap); \
        c->launches += (pin == 0);                                                                              \
    } while (0)
        if (nv <= 1) KMC_REFINE4(1);
        else if (nv <= 2) KMC_REFINE4(2);
        else if (nv <= 4) KMC_REFINE4(4);
        else KMC_REFINE4(8);
#undef KMC_REFINE4
        KMC_CUDA(cudaGetLastError());
        c->launches++;
        // rows whose walk hit the cap: exact distances to all K centroids, register-tiled (none on clustered data)
        return launch_simt_assign(c, overflow, n_over, 0, labels, s, &c->stats()->walk[slot][2]);
    }
#define KMC_REFINE(N)                                                                                       \
    do {                                                                                                    \
        if (sorted)                                                                                         \
            refine_kernel<N, true><<<blocks, 256, 0, s>>>(c->points, (int)c->D, c->centroids(), (int)c->K,  \
                                                         cc, nbr, worklist, nw, labels);                  \
        else                                                                                                \
            refine_kernel<N, false><<<blocks, 256, 0, s>>>(c->points, (int)c->D, c->centroids(), (int)c->K, \
                                                          cc, nbr, worklist, nw, labels);                 \
    } while (0)
    if (npl <= 1) KMC_REFINE(1);
    else if (npl <= 2) KMC_REFINE(2);
    else if (npl <= 4) KMC_REFINE(4);
    else KMC_REFINE(8);
#undef KMC_REFINE
    KMC_CUDA(cudaGetLastError());
    c->launches++;
    return 0;
}

}  // namespace kmc
