// Per-cluster sums, counts and inertia for large K*D, where per-CTA private accumulators do not
// fit on chip and a scatter-add of N*D values through atomics would be ~10x slower than HBM.
//
// FULL pass: the labels are counting-sorted into a row permutation (rows of one cluster become a
// contiguous run) and each run is summed with plain coalesced row loads:
//     histogram -> exclusive scan -> scatter (perm) -> segmented row sum (+ exact inertia, + sum |x|^2)
// One extra streaming pass over the fp32 points, no atomics on the data path.
//
// INCREMENTAL pass (kmc_accumulate from the second call on): Lloyd iterations move fewer and fewer rows, so
// the context keeps the labels of the previous pass and running fp64 sums / counts, finds the rows whose
// label changed (diff_labels_kernel), and runs the same sort + segmented sum over THOSE rows only -- once
// keyed by the new label (+x) and once by the old one (-x).  Every value is added in fp64 (an fp32 point is
// exact in fp64, and the running sums carry 2^-53 relative rounding per addition), so the running sums equal
// a fresh full pass to ~1e-15 whatever the number of iterations.  The inertia of such a pass follows from the
// sums:  sum_i |x_i - c_l(i)|^2 = sum_i |x_i|^2 - 2 sum_k c_k.S_k + sum_k n_k |c_k|^2  (fp64; sum |x|^2 is a
// constant of the data, formed by the first full pass).  Whether a pass is incremental is decided ON THE
// DEVICE (more changed rows than the list holds -> full pass): both variants are enqueued, gated on a flag,
// and the stream never waits for the host.
//
// Replaces reference scripts/distribuitedClustering.py:237-242 (K x equal/where/gather/reduce_mean),
// :244-246 (D2H of the labels + CPU bincount) and :248 (mean*count = sum).  Sums are formed
// directly, so a cluster that is empty on this shard contributes 0, not the reference's NaN.
#include "kmc_common.cuh"
#include "kmc_device.cuh"
#include <atomic>

namespace kmc {

constexpr int kSegRows = 1024;        // rows of one cluster summed by one CTA
constexpr int kScatterChunk = 16384;  // rows per CTA in the scatter pass
// changed rows x D up to which the incremental pass uses atomics.  Measured on one B200: config 3 (8.7 k changed rows x 32)
// 0.121 -> 0.098 ms per accumulation stage with atomics; config 4 (127 k rows x 128 = 16 M values) 0.475 -> 0.53 ms
constexpr int64_t kDeltaAtomicElems = (int64_t)1 << 21;

// Which rows a sort kernel walks: rows r of a contiguous range (rows == nullptr: entry i is row i of `labels`,
// recorded as row_base + i) or the entries of a row list whose length is known on the device only.
struct RowSet {
    const int32_t* labels;                 // contiguous: labels of the range; list: labels of ALL rows (indexed by row)
    const int32_t* rows;                   // list of global row numbers, or nullptr
    int64_t n;                             // entries (contiguous) / capacity of the list
    const unsigned long long* n_dev;       // list: number of valid entries (clamped to n)
    int64_t row_base;
    const unsigned int* gate;              // nullptr, or: run only if *gate == gate_want
    unsigned int gate_want;
    __device__ __forceinline__ bool off() const { return gate && *gate != gate_want; }
    __device__ __forceinline__ int64_t count() const {
        if (!n_dev) return n;
        const unsigned long long v = *n_dev;
        return v < (unsigned long long)n ? (int64_t)v : n;
    }
    __device__ __forceinline__ int label(int64_t i) const { return rows ? labels[rows[i]] : labels[i]; }
    __device__ __forceinline__ int32_t row(int64_t i) const { return rows ? rows[i] : (int32_t)(row_base + i); }
};

// ---- histogram --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hist_kernel(const RowSet rs, int K, unsigned int* __restrict__ hist, int use_smem) {
    extern __shared__ unsigned int lh[];
    if (rs.off()) return;
    const int64_t n = rs.count();
    if (use_smem) {
        for (int i = threadIdx.x; i < K; i += blockDim.x) lh[i] = 0;
        __syncthreads();
    }
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int l = rs.label(i);
        if ((unsigned)l < (unsigned)K) {
            if (use_smem) atomicAdd(&lh[l], 1u); else atomicAdd(&hist[l], 1u);
        }
    }
    if (use_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < K; i += blockDim.x) {
            const unsigned int v = lh[i];
            if (v) atomicAdd(&hist[i], v);
        }
    }
}

// ---- scan: offsets[k] = rows before cluster k; segoff[k] = segments before cluster k ------------
__global__ void __launch_bounds__(1024) scan_kernel(const unsigned int* __restrict__ hist, int K,
                                                    unsigned int* __restrict__ offsets,
                                                    unsigned int* __restrict__ segoff,
                                                    double* __restrict__ counts_out, double sign,
                                                    const unsigned int* __restrict__ gate, unsigned int gate_want) {
    __shared__ unsigned int tot_rows[1024];
    __shared__ unsigned int tot_segs[1024];
    if (gate && *gate != gate_want) return;
    const int t = threadIdx.x;
    const int per = (K + 1023) / 1024;
    const int b = t * per, e = min(K, b + per);
    unsigned int r = 0, sgs = 0;
    for (int i = b; i < e; ++i) {
        const unsigned int h = hist[i];
        r += h;
        sgs += (h + kSegRows - 1) / kSegRows;
    }
    tot_rows[t] = r;
    tot_segs[t] = sgs;
    __syncthreads();
    // Hillis-Steele inclusive scan over the 1024 thread totals
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned int a = 0, c2 = 0;
        if (t >= o) { a = tot_rows[t - o]; c2 = tot_segs[t - o]; }
        __syncthreads();
        tot_rows[t] += a;
        tot_segs[t] += c2;
        __syncthreads();
    }
    unsigned int ro = tot_rows[t] - r, so = tot_segs[t] - sgs;
    for (int i = b; i < e; ++i) {
        const unsigned int h = hist[i];
        offsets[i] = ro;
        segoff[i] = so;
        if (h) counts_out[i] += sign * (double)h;   // chunks of one pass add up (the buffer starts at zero)
        ro += h;
        so += (h + kSegRows - 1) / kSegRows;
    }
    if (t == 1023) { offsets[K] = tot_rows[1023]; segoff[K] = tot_segs[1023]; }
}

// ---- scatter: perm[offsets[l] + rank] = row ----------------------------------------------------
__global__ void __launch_bounds__(256) scatter_kernel(const RowSet rs, int K, const unsigned int* __restrict__ offsets,
                                                      unsigned int* __restrict__ cursor,
                                                      int32_t* __restrict__ perm, int use_smem, int rows_per_cta) {
    extern __shared__ unsigned int sm[];
    if (rs.off()) return;
    const int64_t n = rs.count();
    unsigned int* lcnt = sm;        // [K] local counts, then local cursors
    unsigned int* lbase = sm + K;   // [K] global position of this CTA's first row of each cluster
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r1 = min(n, r0 + (int64_t)rows_per_cta);
    if (r0 >= r1) return;
    if (!use_smem) {
        for (int64_t i = r0 + threadIdx.x; i < r1; i += blockDim.x) {
            const int l = rs.label(i);
            if ((unsigned)l < (unsigned)K) perm[offsets[l] + atomicAdd(&cursor[l], 1u)] = rs.row(i);
        }
        return;
    }
    for (int i = threadIdx.x; i < K; i += blockDim.x) lcnt[i] = 0;
    __syncthreads();
    for (int64_t i = r0 + threadIdx.x; i < r1; i += blockDim.x) {
        const int l = rs.label(i);
        if ((unsigned)l < (unsigned)K) atomicAdd(&lcnt[l], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        const unsigned int cnt = lcnt[i];
        if (cnt) lbase[i] = offsets[i] + atomicAdd(&cursor[i], cnt);
        lcnt[i] = 0;
    }
    __syncthreads();
    for (int64_t i = r0 + threadIdx.x; i < r1; i += blockDim.x) {
        const int l = rs.label(i);
        if ((unsigned)l < (unsigned)K) perm[lbase[l] + atomicAdd(&lcnt[l], 1u)] = rs.row(i);
    }
}

// ---- segmented row sum ---------------------------------------------------------------------------
template <int VEC> struct RowVec;
template <> struct RowVec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float* p) {
        const float4 t = ld_stream_f4(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <> struct RowVec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float* p) { v[0] = ld_stream_f1(p); }
};

// One CTA sums one run of <= kSegRows rows of one cluster.  Each warp owns a contiguous slice of the
// run: it reads 32 permutation entries with one coalesced load, then keeps 8/P rows (512 B each at
// D=128) in flight per lane group before folding them into its accumulators -- the gather is
// latency-bound, so bytes in flight per SM are what sets the bandwidth.  The accumulators are fp64 (the
// running sums of the incremental pass must equal a fresh pass; the DADDs hide under the loads).
// FULL: also the exact difference-form inertia against `cent` and sum |x|^2; otherwise `sign` * sums only.
#ifdef KMC_SEGSUM_MAXNREG
#define KMC_SEGSUM_ATTR __maxnreg__(KMC_SEGSUM_MAXNREG)
#else
#define KMC_SEGSUM_ATTR __launch_bounds__(256)
#endif
template <int VEC, int P, bool FULL>
__global__ void KMC_SEGSUM_ATTR
segsum_kernel(const float* __restrict__ x, int D, int K, const float* __restrict__ cent,
              const int32_t* __restrict__ perm, const unsigned int* __restrict__ offsets,
              const unsigned int* __restrict__ segoff, double* __restrict__ sums, double* __restrict__ scalars,
              int lpr_log2, double sign, const unsigned int* __restrict__ gate, unsigned int gate_want) {
    constexpr int U = 8 / P;                   // row groups in flight
    __shared__ double red[8][32 * VEC * P];
    __shared__ double red_in[8][2];
    if (gate && *gate != gate_want) return;
    const unsigned int seg_total = segoff[K];
    const unsigned int b = blockIdx.x;
    if (b >= seg_total) return;
    // cluster of this segment: largest k with segoff[k] <= b
    int lo = 0, hi = K - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (segoff[mid] <= b) lo = mid; else hi = mid - 1;
    }
    const int k = lo;
    const unsigned int seg = b - segoff[k];
    const unsigned int row_b = offsets[k] + seg * kSegRows;
    const unsigned int row_e = min(offsets[k + 1], row_b + (unsigned)kSegRows);
    const int nrows = (int)(row_e - row_b);

    const int LPR = 1 << lpr_log2;             // lanes cooperating on one row
    const int RPW = 32 >> lpr_log2;            // rows per warp step
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cl = lane & (LPR - 1), sub = lane >> lpr_log2;
    const int col0 = blockIdx.y * (LPR * VEC * P);

    double acc[P][VEC];
    float cen[P][VEC];
    bool on[P];
    int64_t coff[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const int col = col0 + (p * LPR + cl) * VEC;
        on[p] = col < D;
        coff[p] = col;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            acc[p][e] = 0.0;
            cen[p][e] = (FULL && on[p]) ? cent[(int64_t)k * D + col + e] : 0.f;
        }
    }
    double inert = 0.0, xsq = 0.0;
    const int per_warp = (nrows + 7) >> 3;
    const int wb = warp * per_warp;
    const int we = min(nrows, wb + per_warp);
    for (int base = wb; base < we; base += 32) {
        const int cnt = min(32, we - base);
        const int myidx = (lane < cnt) ? perm[row_b + base + lane] : -1;
        const int nsteps = (cnt + RPW - 1) >> (5 - lpr_log2);
        for (int s0 = 0; s0 < nsteps; s0 += U) {
            RowVec<VEC> v[U][P];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = ((s0 + u) << (5 - lpr_log2)) + sub;
                const int idx = __shfl_sync(0xffffffffu, myidx, j & 31);
                ok[u] = (s0 + u < nsteps) && (j < cnt);
                if (ok[u]) {
                    const float* rowp = x + (int64_t)idx * D;
#pragma unroll
                    for (int p = 0; p < P; ++p)
                        if (on[p]) v[u][p].load(rowp + coff[p]);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (ok[u]) {
                    float din = 0.f, dsq = 0.f;          // this lane's share of one row: at most 16 terms in fp32
#pragma unroll
                    for (int p = 0; p < P; ++p)
                        if (on[p]) {
#pragma unroll
                            for (int e = 0; e < VEC; ++e) {
                                const float xv = v[u][p].v[e];
                                acc[p][e] += (double)xv;
                                if (FULL) {
                                    const float t = xv - cen[p][e];
                                    din = fmaf(t, t, din);
                                    dsq = fmaf(xv, xv, dsq);
                                }
                            }
                        }
                    if (FULL) { inert += (double)din; xsq += (double)dsq; }
                }
            }
        }
    }
    // fold the RPW row-groups of a warp together
    for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[p][e] += __shfl_xor_sync(0xffffffffu, acc[p][e], o);
    }
    if (sub == 0) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int e = 0; e < VEC; ++e) red[warp][(p * LPR + cl) * VEC + e] = acc[p][e];
    }
    if (FULL) {
        const double in_w = warp_sum(inert), sq_w = warp_sum(xsq);
        if (lane == 0) { red_in[warp][0] = in_w; red_in[warp][1] = sq_w; }
    }
    __syncthreads();
    const int ncols = LPR * VEC * P;
    for (int cidx = threadIdx.x; cidx < ncols; cidx += blockDim.x) {
        const int col = col0 + cidx;
        if (col < D) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += red[w][cidx];
            atomicAdd(sums + (int64_t)k * D + col, sign * s);
        }
    }
    if (FULL && threadIdx.x < 2) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red_in[w][threadIdx.x];
        atomicAdd(scalars + threadIdx.x, s);       // [0] inertia, [1] sum |x|^2
    }
}

// Sort the rows of `rs` by label into `perm` and add sign * (their per-cluster sums) to sums[K*D], sign * counts
// to counts[K]; full: inertia and sum |x|^2 to scalars[0], scalars[1].  `slot` selects the sort scratch.
static int sorted_accumulate(kmc_ctx* c, const RowSet& rs, int32_t* perm, int slot, double* sums, double* counts,
                             double* scalars, double sign, bool full, cudaStream_t s) {
    const int K = (int)c->K, D = (int)c->D;
    const int64_t n = rs.n;                          // rows (upper bound for a device-counted list)
    if (n <= 0) return 0;
    unsigned int* hist = c->sort_hist(slot);
    unsigned int* cursor = c->sort_cur(slot);
    unsigned int* offsets = c->sort_off(slot);
    unsigned int* segoff = c->sort_seg(slot);
    // hist and cursor are adjacent in the workspace: one memset clears both
    KMC_CUDA(cudaMemsetAsync(hist, 0, (size_t)c->sort_offsets, s));

    const int hist_smem_ok = (size_t)K * 4 <= 96 * 1024;
    {
        int64_t blocks = (n + 256 * 16 - 1) / (256 * 16);
        const int64_t cap = (int64_t)c->sm_count * 8;
        blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
        const size_t sm = hist_smem_ok ? (size_t)K * 4 : 0;
        if (sm > 48 * 1024)
            KMC_CUDA(cudaFuncSetAttribute(hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        hist_kernel<<<(unsigned)blocks, 256, sm, s>>>(rs, K, hist, hist_smem_ok);
        KMC_CUDA(cudaGetLastError());
    }
    scan_kernel<<<1, 1024, 0, s>>>(hist, K, offsets, segoff, counts, sign, rs.gate, rs.gate_want);
    KMC_CUDA(cudaGetLastError());
    {
        const int sc_smem_ok = (size_t)K * 8 <= 192 * 1024;
        const size_t sm = sc_smem_ok ? (size_t)K * 8 : 0;
        // a row list is short and its labels are gathered (latency-bound): fewer rows per CTA, more CTAs at work
        const int per_cta = rs.rows ? (K <= 2048 ? 2048 : kScatterChunk / 2) : kScatterChunk;
        const int64_t blocks = (n + per_cta - 1) / per_cta;
        if (sm > 48 * 1024)
            KMC_CUDA(cudaFuncSetAttribute(scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        scatter_kernel<<<(unsigned)blocks, 256, sm, s>>>(rs, K, offsets, cursor, perm, sc_smem_ok, per_cta);
        KMC_CUDA(cudaGetLastError());
    }
    {
        const bool vec4 = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(c->points) & 15) == 0);
        const int vec = vec4 ? 4 : 1;
        int lpr_log2 = 0;
        while ((1 << lpr_log2) < 32 && (1 << lpr_log2) * vec < D) ++lpr_log2;
        const int cols_per_pass = (1 << lpr_log2) * vec;
        int passes = (D + cols_per_pass - 1) / cols_per_pass;
        passes = passes >= 4 ? 4 : (passes >= 2 ? 2 : 1);
        const int cols_per_block = cols_per_pass * passes;
        const unsigned gy = (unsigned)((D + cols_per_block - 1) / cols_per_block);
        const int64_t gx = (n + kSegRows - 1) / kSegRows + K;
        dim3 grid((unsigned)gx, gy);
#define KMC_SEGSUM2(V, PP, F)                                                                                   \
    segsum_kernel<V, PP, F><<<grid, 256, 0, s>>>(c->points, D, K, c->centroids(), perm, offsets, segoff, sums, \
                                                 scalars, lpr_log2, sign, rs.gate, rs.gate_want)
#define KMC_SEGSUM(V, PP) do { if (full) KMC_SEGSUM2(V, PP, true); else KMC_SEGSUM2(V, PP, false); } while (0)
        if (vec4) {
            if (passes == 1) KMC_SEGSUM(4, 1); else if (passes == 2) KMC_SEGSUM(4, 2); else KMC_SEGSUM(4, 4);
        } else {
            if (passes == 1) KMC_SEGSUM(1, 1); else if (passes == 2) KMC_SEGSUM(1, 2); else KMC_SEGSUM(1, 4);
        }
#undef KMC_SEGSUM
#undef KMC_SEGSUM2
        KMC_CUDA(cudaGetLastError());
    }
    c->launches += 4;
    return 0;
}

// sums / counts / inertia / sum |x|^2 of rows [r0, r1), ADDED to the partial buffer; `gate`: see RowSet
int launch_sorted_accumulate(kmc_ctx* c, const int32_t* labels_all, int64_t r0, int64_t r1, int chunk,
                             const unsigned int* gate, cudaStream_t s) {
    RowSet rs;
    rs.labels = labels_all + r0; rs.rows = nullptr; rs.n = r1 - r0; rs.n_dev = nullptr; rs.row_base = r0;
    rs.gate = gate; rs.gate_want = 1u;
    double* part = c->partials();
    return sorted_accumulate(c, rs, c->perm() + r0, chunk, part, part + c->K * c->D, part + c->K * c->D + c->K, 1.0, true, s);
}

// ---- incremental pass ----------------------------------------------------------------------------
// Rows whose label differs from the previous pass, in ascending order within a warp's 32 rows; the list holds
// `cap` rows, the count keeps running beyond it (the gate below then asks for a full pass).
__global__ void __launch_bounds__(256) diff_labels_kernel(const int32_t* __restrict__ cur, const int32_t* __restrict__ prev,
                                                          int64_t n, int32_t* __restrict__ list, unsigned long long cap,
                                                          unsigned long long* __restrict__ n_chg) {
    // a warp compares 4 x 128 consecutive rows per step: four 16-byte loads per lane and label array in flight
    // (both label buffers are 256-byte aligned workspace regions)
    constexpr int U = 4;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t base = warp * (128 * U); base < n; base += nwarps * (128 * U)) {
        int4 a[U], b[U];
        unsigned bits[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i0 = base + 128 * u + 4 * lane;
            if (i0 + 3 < n) {
                a[u] = __ldcs(reinterpret_cast<const int4*>(cur + i0));
                b[u] = __ldcs(reinterpret_cast<const int4*>(prev + i0));
            }
        }
        bool any = false;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i0 = base + 128 * u + 4 * lane;
            bits[u] = 0;
            if (i0 + 3 < n) {
                bits[u] = (a[u].x != b[u].x ? 1u : 0u) | (a[u].y != b[u].y ? 2u : 0u) | (a[u].z != b[u].z ? 4u : 0u) |
                          (a[u].w != b[u].w ? 8u : 0u);
            } else {
                for (int e = 0; e < 4; ++e)
                    if (i0 + e < n && cur[i0 + e] != prev[i0 + e]) bits[u] |= 1u << e;
            }
            any |= bits[u] != 0u;
        }
        if (__ballot_sync(0xffffffffu, any) == 0u) continue;
        // one slot reservation (a returning atomic: ~1 us) per warp step, not one per 128 rows: list order within the
        // step is lane-major, which keeps the rows of a lane (16 consecutive ones) together
        int cnt = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) cnt += __popc(bits[u]);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        unsigned long long pos = 0;
        if (lane == 0) pos = atomicAdd(n_chg, (unsigned long long)total);
        pos = __shfl_sync(0xffffffffu, pos, 0) + (unsigned long long)(incl - cnt);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i0 = base + 128 * u + 4 * lane;
            unsigned w = bits[u];
            while (w) {
                const int e = __ffs(w) - 1;
                w &= w - 1u;
                if (pos < cap) list[pos] = (int32_t)(i0 + e);
                ++pos;
            }
        }
    }
}

// acc_full = 1 and an empty list if more rows changed than the list holds; otherwise the changed rows are applied to
// the running sums by the sort + segmented sum (acc_sorted = 1) or, up to `atomic_rows` of them, with fp64 atomics
__global__ void delta_gate_kernel(DeviceStats* st, unsigned long long cap, unsigned long long atomic_rows) {
    const unsigned long long m = st->n_changed;
    if (m > cap) { st->acc_full = 1u; st->acc_sorted = 0u; st->n_changed_eff = 0ull; }
    else { st->acc_full = 0u; st->acc_sorted = (m > atomic_rows) ? 1u : 0u; st->n_changed_eff = m; }
}

// Few changed rows (the steady state of a Lloyd run: 0.1 % of the rows per iteration at config 4): +x under the new
// label and -x under the old one straight into the running fp64 sums, one warp per row.  The sort + segmented sum
// needs ten launches for the same effect and only pays when the atomics of one cluster start to queue up.
__global__ void __launch_bounds__(256) delta_atomic_kernel(const float* __restrict__ x, int D, int K,
                                                           const int32_t* __restrict__ cur, const int32_t* __restrict__ prev,
                                                           const int32_t* __restrict__ list, double* __restrict__ run,
                                                           const DeviceStats* __restrict__ st) {
    if (st->acc_full != 0u || st->acc_sorted != 0u) return;
    const int64_t m = (int64_t)st->n_changed_eff;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    double* counts = run + (int64_t)K * D;
    for (int64_t i = warp; i < m; i += nwarps) {
        const int64_t row = list[i];
        const int ln = cur[row], lo = prev[row];
        const bool okn = (unsigned)ln < (unsigned)K, oko = (unsigned)lo < (unsigned)K;
        const float* xr = x + row * D;
        for (int j = lane; j < D; j += 32) {
            const double v = (double)ld_stream_f1(xr + j);
            if (okn) atomicAdd(run + (int64_t)ln * D + j, v);
            if (oko) atomicAdd(run + (int64_t)lo * D + j, -v);
        }
        if (lane == 0) {
            if (okn) atomicAdd(counts + ln, 1.0);
            if (oko) atomicAdd(counts + lo, -1.0);
        }
    }
}

// After the two variants: a full pass leaves fresh sums in `part` -> they become the running state; an
// incremental pass updated the running state -> it is copied into `part`, and the inertia follows from the sums.
// Layout of both buffers: [K*D sums | K counts | inertia | sum |x|^2].
__global__ void __launch_bounds__(256) delta_finish_kernel(double* __restrict__ part, double* __restrict__ run,
                                                           const float* __restrict__ cent, int K, int D,
                                                           const DeviceStats* __restrict__ st) {
    __shared__ double red[8];
    const int64_t KD = (int64_t)K * D, total = KD + K;
    const bool full = st->acc_full != 0u;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        if (full) { run[i] = part[i]; continue; }
        const double v = run[i];
        part[i] = v;
        if (i < KD) {
            const double cnt = run[KD + i / D];
            if (cnt > 0.0) {                                  // (an empty cluster may hold a NaN centroid)
                const double cv = (double)cent[i];
                acc += cv * (cnt * cv - 2.0 * v);
            }
        }
    }
    if (full) {
        if (blockIdx.x == 0 && threadIdx.x == 0) run[total + 1] = part[total + 1];
        return;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[w];
        if (blockIdx.x == 0) { s += run[total + 1]; part[total + 1] = run[total + 1]; }
        atomicAdd(part + total, s);
    }
}

// The accumulation stage of kmc_accumulate for the paths that sort (SIMT, tensor): labels of this pass are in
// c->labels(), those of the previous pass in c->labels_prev().  Leaves the partial buffer filled and swaps the
// two label buffers.
int launch_accumulate_stage(kmc_ctx* c, cudaStream_t s) {
    const int64_t KD = c->K * c->D;
    double* part = c->partials();
    double* run = c->running();
    DeviceStats* st = c->stats();
    if (!c->delta_enabled || !c->delta_valid) {
        KMC_TRY(launch_sorted_accumulate(c, c->labels(), 0, c->n, 0, nullptr, s));
        if (c->delta_enabled) {
            KMC_CUDA(cudaMemcpyAsync(run, part, (size_t)(KD + c->K + 2) * 8, cudaMemcpyDeviceToDevice, s));
            c->delta_valid = true;
        }
        KMC_CUDA(cudaMemsetAsync(&st->n_changed, 0, sizeof(unsigned long long), s));
        c->swap_labels();
        return 0;
    }
    const unsigned long long cap = (unsigned long long)c->chg_cap;
    KMC_CUDA(cudaMemsetAsync(&st->n_changed, 0, sizeof(unsigned long long), s));
    {
        int64_t blocks = (c->n + 256 * 16 - 1) / (256 * 16);
        const int64_t bcap = (int64_t)c->sm_count * 8;
        blocks = blocks > bcap ? bcap : (blocks < 1 ? 1 : blocks);
        diff_labels_kernel<<<(unsigned)blocks, 256, 0, s>>>(c->labels(), c->labels_prev(), c->n, c->changed(), cap, &st->n_changed);
        KMC_CUDA(cudaGetLastError());
        // default: atomics up to 2 M values' worth of changed rows (kDeltaAtomicElems)
        const int64_t atomic_rows = c->delta_atomic_rows >= 0 ? c->delta_atomic_rows : kDeltaAtomicElems / (c->D > 0 ? c->D : 1);
        delta_gate_kernel<<<1, 1, 0, s>>>(st, cap, (unsigned long long)atomic_rows);
        KMC_CUDA(cudaGetLastError());
        c->launches += 2;
        if (atomic_rows > 0 && c->chg_cap > 0) {
            delta_atomic_kernel<<<c->sm_count * 8, 256, 0, s>>>(c->points, (int)c->D, (int)c->K, c->labels(), c->labels_prev(),
                                                                c->changed(), run, st);
            KMC_CUDA(cudaGetLastError());
            c->launches++;
        }
    }
    if (c->chg_cap > 0) {
        RowSet rs;
        rs.rows = c->changed(); rs.n = c->chg_cap; rs.n_dev = &st->n_changed_eff; rs.row_base = 0;
        rs.gate = &st->acc_sorted; rs.gate_want = 1u;
        rs.labels = c->labels();           // + x under the new label
        KMC_TRY(sorted_accumulate(c, rs, c->perm(), kMaxChunks, run, run + KD, nullptr, 1.0, false, s));
        rs.labels = c->labels_prev();      // - x under the old one
        KMC_TRY(sorted_accumulate(c, rs, c->perm(), kMaxChunks, run, run + KD, nullptr, -1.0, false, s));
    }
    KMC_TRY(launch_sorted_accumulate(c, c->labels(), 0, c->n, 0, &st->acc_full, s));
    {
        int64_t blocks = (KD + c->K + 255) / 256;
        const int64_t bcap = (int64_t)c->sm_count * 4;
        blocks = blocks > bcap ? bcap : blocks;
        delta_finish_kernel<<<(unsigned)blocks, 256, 0, s>>>(part, run, c->centroids(), (int)c->K, (int)c->D, st);
        KMC_CUDA(cudaGetLastError());
        c->launches++;
    }
    c->swap_labels();
    return 0;
}

}  // namespace kmc

// =================================================================================================
// Allocation-free entry point (kmc_cluster_sums_noalloc, see include/kmc.h)
//
// Some hosts may not allocate device memory at all: an operator that a framework calls with only its
// input and output buffers (the framework owns every byte of HBM through its own pool and hands the
// operator no scratch), or a serving loop with a fixed memory budget where a cudaMalloc in the middle
// of a request would synchronise the whole device or fail outright.  Those callers cannot follow the
// kmc_workspace_bytes / kmc_bind_workspace protocol.  For them the scratch of the sorted accumulation
// (histogram, cursors, offsets, segment offsets and the row permutation) lives in one file-scope
// __device__ array of a fixed, modest size.  Every GPU that loads this library carries the array
// until the process exits, every call on a device shares it, and it keeps whatever the previous call
// left in it.
// =================================================================================================
namespace kmc {

constexpr int64_t kNoAllocMaxRows = 1 << 20;     // rows per call
constexpr int64_t kNoAllocMaxK = 4096;           // clusters per call
constexpr int kNoAllocMaxDevices = 64;

struct NoAllocScratch {
    unsigned int hist[kNoAllocMaxK];             // hist and cursor are cleared together at the start of a call
    unsigned int cursor[kNoAllocMaxK];
    unsigned int offsets[kNoAllocMaxK + 64];
    unsigned int segoff[kNoAllocMaxK + 64];
    int32_t perm[kNoAllocMaxRows];
};

// In host code this name is only a handle for the runtime (cudaGetSymbolAddress below), never an address
// a kernel could dereference; the kernels get the device address looked up once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

// Per device: the address of the scratch array and the SM count, published together.  Two host threads may make
// the first call on a device at the same time: the pointer is stored last, with release ordering, and read with
// acquire ordering, so a thread that sees it also sees the SM count (both threads would store the same values).
static std::atomic<NoAllocScratch*> g_noalloc_ptr[kNoAllocMaxDevices];
static int g_noalloc_sms[kNoAllocMaxDevices] = {};

// First call on a device: check that it is an sm_100 GPU (this library holds no other code) and resolve the
// device address of the scratch array, with that device current.
static int noalloc_setup(int device) {
    if (g_noalloc_ptr[device].load(std::memory_order_acquire)) return 0;
    int major = 0, sms = 0;
    KMC_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10)
        KMC_FAIL(KMC_E_DEVICE, "device %d is sm_%dx; this library contains sm_100a code only (no fallback)", device, major);
    void* ptr = nullptr;
    KMC_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    KMC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    g_noalloc_sms[device] = sms;
    g_noalloc_ptr[device].store(static_cast<NoAllocScratch*>(ptr), std::memory_order_release);
    return 0;
}

// the calling thread's current device is put back when the entry point returns
struct DeviceRestore {
    int prev = -1;
    ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace kmc

extern "C" int kmc_cluster_sums_noalloc(int device, const float* dev_points, const int32_t* dev_labels,
                                        const float* dev_centroids, int64_t n_rows, int64_t D, int64_t K,
                                        double* dev_partials, void* stream) {
    using namespace kmc;
    if (device < 0 || device >= kNoAllocMaxDevices) KMC_FAIL(KMC_E_INVALID, "device index %d out of range", device);
    if (D < 1 || K < 1 || n_rows < 0) KMC_FAIL(KMC_E_INVALID, "need n_rows >= 0, D >= 1, K >= 1");
    if (n_rows > kNoAllocMaxRows)
        KMC_FAIL(KMC_E_NOMEM, "kmc_cluster_sums_noalloc handles at most %lld rows per call (got %lld): its scratch is "
                 "a fixed-size device array; use a kmc_ctx with a caller-provided workspace for larger shards",
                 (long long)kNoAllocMaxRows, (long long)n_rows);
    if (K > kNoAllocMaxK)
        KMC_FAIL(KMC_E_NOMEM, "kmc_cluster_sums_noalloc handles at most %lld clusters (got %lld)",
                 (long long)kNoAllocMaxK, (long long)K);
    if (K * D >= ((int64_t)1 << 31)) KMC_FAIL(KMC_E_INVALID, "K*D too large");
    if (!dev_partials || (n_rows > 0 && (!dev_points || !dev_labels || !dev_centroids)))
        KMC_FAIL(KMC_E_INVALID, "null argument");
    DeviceRestore restore;
    if (cudaGetDevice(&restore.prev) != cudaSuccess) restore.prev = -1;
    KMC_CUDA(cudaSetDevice(device));
    KMC_TRY(noalloc_setup(device));
    NoAllocScratch* sc = g_noalloc_ptr[device].load(std::memory_order_acquire);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int Ki = (int)K, Di = (int)D;

    // the output accumulates through atomics and the histogram / cursors count up: all start at zero
    KMC_CUDA(cudaMemsetAsync(dev_partials, 0, (size_t)(K * D + K + 2) * sizeof(double), s));
    KMC_CUDA(cudaMemsetAsync(sc->hist, 0, 2 * sizeof(unsigned int) * (size_t)kNoAllocMaxK, s));
    if (n_rows == 0) return 0;

    RowSet rs;
    rs.labels = dev_labels; rs.rows = nullptr; rs.n = n_rows; rs.n_dev = nullptr; rs.row_base = 0;
    rs.gate = nullptr; rs.gate_want = 0u;
    {
        int64_t blocks = (n_rows + 256 * 16 - 1) / (256 * 16);
        const int64_t cap = (int64_t)g_noalloc_sms[device] * 8;
        blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
        hist_kernel<<<(unsigned)blocks, 256, (size_t)Ki * 4, s>>>(rs, Ki, sc->hist, 1);
        KMC_CUDA(cudaGetLastError());
    }
    scan_kernel<<<1, 1024, 0, s>>>(sc->hist, Ki, sc->offsets, sc->segoff, dev_partials + K * D, 1.0, nullptr, 0u);
    KMC_CUDA(cudaGetLastError());
    {
        const int64_t blocks = (n_rows + kScatterChunk - 1) / kScatterChunk;
        scatter_kernel<<<(unsigned)blocks, 256, (size_t)Ki * 8, s>>>(rs, Ki, sc->offsets, sc->cursor, sc->perm, 1, kScatterChunk);
        KMC_CUDA(cudaGetLastError());
    }
    {
        const bool vec4 = (Di % 4 == 0) && ((reinterpret_cast<uintptr_t>(dev_points) & 15) == 0);
        const int vec = vec4 ? 4 : 1;
        int lpr_log2 = 0;
        while ((1 << lpr_log2) < 32 && (1 << lpr_log2) * vec < Di) ++lpr_log2;
        const int cols_per_pass = (1 << lpr_log2) * vec;
        int passes = (Di + cols_per_pass - 1) / cols_per_pass;
        passes = passes >= 4 ? 4 : (passes >= 2 ? 2 : 1);
        const int cols_per_block = cols_per_pass * passes;
        dim3 grid((unsigned)((n_rows + kSegRows - 1) / kSegRows + K), (unsigned)((Di + cols_per_block - 1) / cols_per_block));
#define KMC_SEGSUM_NA(V, PP)                                                                                 \
    segsum_kernel<V, PP, true><<<grid, 256, 0, s>>>(dev_points, Di, Ki, dev_centroids, sc->perm, sc->offsets, sc->segoff, \
                                                    dev_partials, dev_partials + K * D + K, lpr_log2, 1.0, nullptr, 0u)
        if (vec4) {
            if (passes == 1) KMC_SEGSUM_NA(4, 1); else if (passes == 2) KMC_SEGSUM_NA(4, 2); else KMC_SEGSUM_NA(4, 4);
        } else {
            if (passes == 1) KMC_SEGSUM_NA(1, 1); else if (passes == 2) KMC_SEGSUM_NA(1, 2); else KMC_SEGSUM_NA(1, 4);
        }
#undef KMC_SEGSUM_NA
        KMC_CUDA(cudaGetLastError());
    }
    return 0;
}
