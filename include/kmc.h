/*
 * kmc.h -- C ABI of libkmc_b200.so: the data-parallel K-Means Lloyd iteration of
 * Jhonsonzhangxing/tensorflow-distributed-clustering, hand-written for NVIDIA B200 (sm_100a).
 *
 * The reference has no FFI / op-registration layer: its boundary is the Python function
 * `distribuited_k_means` (reference scripts/distribuitedClustering.py:180-294).  This header is
 * the boundary that a Python (ctypes), C or C++ host binds instead of building that TensorFlow
 * graph.  Each entry point cites the reference lines whose work it replaces.
 *
 * Conventions
 *   - One `kmc_ctx` drives ONE shard of rows on ONE GPU (one process per GPU, or one ctx per
 *     device inside a single process).  Cross-GPU reduction of the partial buffer is the host's
 *     job (NCCL all-reduce of `kmc_partials_len` doubles) between kmc_accumulate and kmc_update.
 *   - All pointers named `dev_*` are device pointers valid on the ctx's device; `host_*` are
 *     host pointers.  `stream` is a `cudaStream_t` passed as `void*` (NULL = default stream).
 *     Every call is asynchronous on `stream` unless it returns values to the host.
 *   - The caller provides the device memory that scales with the shard: one workspace
 *     (`kmc_workspace_bytes` / `kmc_bind_workspace`) and the points buffer, and keeps both alive
 *     until `kmc_destroy`.  The library itself allocates only fixed-size staging: two pinned host buffers and,
 *     for fp64 sources / int64 label delivery, two device buffers of 64 MB each on first use of
 *     kmc_stage_points / kmc_labels_to_host (freed by kmc_destroy).
 *   - Return value: 0 = OK; > 0 = a cudaError_t; < 0 = a library error (KMC_E_*).  A
 *     human-readable message for the last failure on the calling thread is in kmc_last_error().
 *   - There is no CPU fallback: every entry point runs CUDA kernels or fails.
 */
#ifndef KMC_B200_H_
#define KMC_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kmc_ctx kmc_ctx;

/* library error codes (negative) */
#define KMC_E_INVALID   (-1)   /* bad argument / shape */
#define KMC_E_STATE     (-2)   /* call order (workspace / points / centroids not bound yet) */
#define KMC_E_NOMEM     (-3)   /* workspace too small */
#define KMC_E_UNSUPPORTED (-4) /* requested kernel path cannot run this shape */
#define KMC_E_DEVICE    (-5)   /* device is not an sm_100 GPU / driver entry point missing */

/* kmc_create flags */
#define KMC_EMPTY_KEEP      0x0u  /* globally empty cluster keeps its previous centroid (default) */
#define KMC_EMPTY_NAN       0x1u  /* ... becomes NaN, as the reference's 0/0 at :261 */
#define KMC_PATH_AUTO       0x00u /* pick the kernel path from (K, D) */
#define KMC_PATH_SMALL      0x10u /* fused FFMA assign+accumulate, centroids in shared memory.  For K <= 64 the
                                     argmin compares distances whose lowest ceil(log2 K) mantissa bits carry the
                                     centroid index: two distances closer than 2^-(23-bits) relative (<= 2^-17, inside
                                     the 1e-5 near-tie window) resolve to the lower index, and the inertia sums the
                                     truncated distances (low by < 2^-(23-bits) relative).  Sums, counts and the
                                     inertia are added in a fixed order: results are bit-reproducible for a shard. */
#define KMC_PATH_SIMT       0x20u /* exact fp32 register-tiled assign + sorted accumulate */
#define KMC_PATH_TC         0x30u /* tcgen05/TMEM/TMA split-fp16 assign + exact refine + sorted accumulate */
#define KMC_PATH_MASK       0xF0u

/* kmc_stage_points source dtypes */
#define KMC_DTYPE_F32 0
#define KMC_DTYPE_F64 1

/* index of values returned by kmc_read_stats */
#define KMC_STAT_INERTIA     0  /* sum_i min_k |x_i - c_k|^2 of the last accumulate (this shard) */
#define KMC_STAT_SHIFT2      1  /* sum_k |c_new - c_old|^2 of the last update */
#define KMC_STAT_AMBIGUOUS   2  /* rows the tensor path sent to the exact refine kernel, last pass */
#define KMC_STAT_KERNEL_PATH 3  /* KMC_PATH_* actually used */
#define KMC_STAT_LAUNCHES    4  /* kernels launched by this ctx since creation */
#define KMC_STAT_X_SCALED_MAX 5 /* tensor path: max |x| times the scale of the fp16 point planes; must stay
                                  below 32768 (only a streamed load can exceed it, see kmc_accumulate_rows) */
#define KMC_STAT_CHANGED     6  /* rows whose label changed in the last kmc_accumulate (all rows for a full pass
                                  decided by the host: first pass, after a rebind) */
#define KMC_STAT_FULL_PASS   7  /* 1: the last kmc_accumulate summed every row; 0: it updated the running sums
                                  with the changed rows only (KMC_PARAM_INCREMENTAL) */
#define KMC_STAT_WALK_OVERFLOW 8 /* tensor path: refined rows of the last pass whose candidate walk hit its cap and
                                  that the register-tiled exact kernel decided instead */
#define KMC_STAT_COUNT       12

/* Version of this ABI (major*1000 + minor). */
int kmc_version(void);

/* Message describing the last error on the calling thread ("" if none). */
const char* kmc_last_error(void);

/*
 * Create the context for one shard: `n_rows` points of dimension `D`, `K` clusters, on CUDA
 * device `device`.  Replaces the per-GPU graph construction of reference :201-251
 * (`with tf.device(GPU_name)` ... ) -- here nothing is traced; the call only picks the kernel
 * path and lays out the workspace.
 */
int kmc_create(kmc_ctx** out, int device, int64_t n_rows, int64_t D, int64_t K, uint32_t flags);

/* Free everything the library owns (streams/events/pinned staging).  The workspace and the
 * points stay the caller's.  Reference: end of the `with tf.Session` block, :271-282. */
void kmc_destroy(kmc_ctx* ctx);

/* Bytes of device workspace this ctx needs (labels, sort permutation, partial sums, split
 * fp16 planes of the points for the tensor path, ...). */
int64_t kmc_workspace_bytes(const kmc_ctx* ctx);

/* Hand the workspace to the ctx (256-byte aligned device pointer). */
int kmc_bind_workspace(kmc_ctx* ctx, void* dev_workspace, int64_t bytes);

/*
 * Bind the shard's points: row-major fp32 [n_rows, D], resident on the device for all
 * iterations.  Replaces `X = tf.Variable(X_mat)` (:217) + the initializer run (:273).  On the
 * tensor path this also launches the one-time split of the points into the scaled fp16
 * hi/lo planes the tcgen05 kernel reads.
 */
int kmc_bind_points(kmc_ctx* ctx, const float* dev_points, void* stream);

/*
 * Streamed loading (tensor path): overlap the first Lloyd iteration with the H2D copy of the points.
 * kmc_bind_points_streamed binds the (still empty) device buffer; as soon as rows [r0, r1) have arrived
 * kmc_accumulate_rows splits them into the fp16 planes, assigns them and ADDS their sums / counts / inertia
 * to the partial buffer (zeroed by the piece that starts at row 0).  Pieces are consecutive, start at row 0 and
 * end at multiples of 128 (the last one at n_rows); after the last piece the context is in the same state as
 * after kmc_bind_points + kmc_accumulate, so kmc_update and further kmc_accumulate / kmc_iterate calls follow.
 * The plane scale is fixed by the first piece: if KMC_STAT_X_SCALED_MAX reads >= 32768 afterwards (later
 * rows more than 8x larger than every row of the first piece) rebind with kmc_bind_points and repeat the
 * iteration.  KMC_E_UNSUPPORTED on the small-D and SIMT paths (bind normally there).  No counterpart in the
 * reference, whose feed of `X` (:273) completes before the first `sess.run` of the loop (:277).
 */
int kmc_bind_points_streamed(kmc_ctx* ctx, const float* dev_points, void* stream);
int kmc_accumulate_rows(kmc_ctx* ctx, int64_t r0, int64_t r1, void* stream);

/*
 * Pinned, double-buffered host->device staging of rows [row0, row0+n_rows) of a host array
 * (fp32 or fp64, row-major [*, D]) into `dev_dst` (fp32 [n_rows, D]); pageable fp64 rows are converted
 * to fp32 by the host threads that copy them into the pinned chunks, pinned fp64 rows on the device
 * (same rounding either way).  Replaces the feed of `all_data` + `tf.split` + per-GPU variable init (:196-197, :217,
 * :273).  `dev_dst` is any fp32 device buffer (normally a slice of the points buffer).
 * Returns after the last chunk has been queued and the host buffer may be reused.
 */
int kmc_stage_points(kmc_ctx* ctx, float* dev_dst, const void* host_src, int src_dtype,
                     int64_t row0, int64_t n_rows, void* stream);

/* Set / read the K x D fp32 centroids (host, row-major).  Replaces `global_centroids =
 * tf.Variable(initial_centers)` (:199) and the fetch of `global_centroids` (:279). */
int kmc_set_centroids(kmc_ctx* ctx, const float* host_centroids, void* stream);
int kmc_get_centroids(kmc_ctx* ctx, float* host_centroids, void* stream);

/* Device address of the centroids (fp32 [K, D]) inside the workspace, for hosts that want to
 * broadcast or inspect them without a host round trip. */
float* kmc_centroids_ptr(kmc_ctx* ctx);

/*
 * One assignment + accumulation pass over the shard against the current centroids: nearest
 * centroid per row (lowest index on ties), per-cluster sums, counts and the inertia.
 * Replaces the distance / argmin sub-graph (:221-234), the K per-cluster gather+mean ops
 * (:237-242), the CPU bincount (:244-246) and the mean*count reconstruction (:248).
 * The N x K distance matrix is never formed.  Result: the partial buffer
 *   double[K*D] sums (row-major) | double[K] counts | double[1] inertia | double[1] sum_i |x_i|^2
 * at kmc_partials_ptr(); it is overwritten by every call.
 *
 * Incremental accumulation (SIMT and tensor paths, KMC_PARAM_INCREMENTAL, on by default): the context keeps the
 * labels of its previous pass and running fp64 sums / counts.  From the second call on, only the rows whose
 * label changed are summed (added under the new label, subtracted under the old one, every value in fp64, so
 * the running sums equal a fresh pass to ~1e-15), and the inertia follows from the sums as
 * sum|x|^2 - 2 sum_k c_k.S_k + sum_k n_k|c_k|^2.  The decision is taken on the device: when more than a quarter
 * of the rows changed, the same call sums every row as the first one does.  All N x K distances are evaluated
 * in either case; results are those of the full pass.  Rebinding points or workspace, kmc_resize and
 * kmc_accumulate_add reset the running state.
 */
int kmc_accumulate(kmc_ctx* ctx, void* stream);
double* kmc_partials_ptr(kmc_ctx* ctx);
int64_t kmc_partials_len(const kmc_ctx* ctx);   /* = K*D + K + 2 */

/*
 * Out-of-core batches (points that do not fit in HBM at once): kmc_resize changes the number of rows
 * of the NEXT batch (at most the n_rows of kmc_create; bind its points afterwards), and
 * kmc_accumulate_add works like kmc_accumulate but ADDS the batch to the partial buffer instead of
 * overwriting it, so several batches add up to one exact Lloyd iteration before kmc_update.
 * Replaces the reference's batch loop (main :328-360, run_experiments :296-318), which instead runs
 * an independent K-Means per batch and averages the end centres (:310).
 */
int kmc_resize(kmc_ctx* ctx, int64_t n_rows);
int kmc_accumulate_add(kmc_ctx* ctx, void* stream);

/*
 * Centroid update from the (all-reduced) partial buffer: c[k] = sums[k] / counts[k], the squared
 * centre shift, and on the tensor path the next iteration's split centroid planes and norms.
 * Replaces add_n / tile / div / transpose / assign on /cpu:0 (:257-263).
 */
int kmc_update(kmc_ctx* ctx, void* stream);

/*
 * Single-GPU Lloyd loop: up to `max_iters` x (kmc_accumulate, kmc_update) on `stream`, stopping
 * early when shift2 <= tol (tol <= 0: never, the reference's behaviour).  Replaces the
 * `for i in range(n_max_iters): sess.run(...)` loop (:277-280).  Synchronises the stream before
 * returning.  Outputs (any may be NULL): iterations done, inertia of the last accumulate,
 * shift2 of the last update.
 */
int kmc_iterate(kmc_ctx* ctx, int max_iters, double tol, int* n_iter_out,
                double* inertia_out, double* shift2_out, void* stream);

/*
 * Label-only pass against the current centroids: dev_labels[i] = argmin_k |x_i - c_k|^2
 * (int32, n_rows entries).  Replaces the second `sess.run(result_matrix, ...)` that ships the
 * whole N x K matrix to the CPU for an argmin (:235, :255, :282).
 */
int kmc_assign(kmc_ctx* ctx, int32_t* dev_labels, void* stream);

/*
 * Labels of the last kmc_accumulate pass (int32 [n_rows] into dev_labels): argmin against the centroids that
 * pass used, i.e. BEFORE the kmc_update that followed it.  This is what the reference fetches as
 * `best_centroids` in the same sess.run as the update (:279) and returns under the key `centroids` (:286).
 * SIMT / tensor paths: a device copy of the labels the pass kept; small-D path (which keeps none): one
 * label-only pass against the saved pre-update centroids.
 */
int kmc_last_labels(kmc_ctx* ctx, int32_t* dev_labels, void* stream);

/*
 * Deliver `n_labels` int32 device labels (as written by kmc_assign) to HOST memory as int32
 * (label_bytes = 4) or int64 (label_bytes = 8: the dtype of the reference's `np.argmin(...)` result,
 * :282; widened on the device).  `host_labels` may be pageable -- chunks are staged through the
 * context's pinned double buffer and copied out by host threads while the next chunk is in flight -- or
 * pinned (direct DMA).  Synchronises the stream.  Together with kmc_assign this replaces the reference's
 * D2H of the whole N x K float64 distance matrix (:235, :255, :282).
 */
int kmc_labels_to_host(kmc_ctx* ctx, const int32_t* dev_labels, void* host_labels, int label_bytes,
                       int64_t n_labels, void* stream);

/*
 * dev_out[i, :] = centroids[dev_labels[i], :] (fp32 [n_rows, D] on the device; NaN rows for labels outside
 * [0, K)): the recoloured points of the colour-quantisation use case, which the reference's notebook forms on
 * the host as `center[label.flatten()]` (notebooks/Testing Images.ipynb:1233-1235).
 */
int kmc_gather_centroids(kmc_ctx* ctx, const int32_t* dev_labels, float* dev_out, void* stream);

/* Synchronise `stream` and copy KMC_STAT_COUNT doubles (see KMC_STAT_*) to the host. */
int kmc_read_stats(kmc_ctx* ctx, double* host_stats, void* stream);

/* Tunable parameters (advanced; defaults are chosen at kmc_create). */
#define KMC_PARAM_TC_GAMMA  1   /* relative error bound of the split-fp16 tensor scores used to
                                   decide which rows go to the exact refine kernel */
#define KMC_PARAM_TC_TERMS  3   /* tensor path: MMA terms per score. 1 = hi*hi only (fp32 accumulators) + exact refine
                                   of every uncertified row; 16 = the same term with fp16 accumulators (packed TMEM
                                   reads, half2 epilogue: the cheapest pass, the widest certification margin);
                                   3 = fp32-accurate split; 0 (default) = start with 16 (K > 256, see
                                   KMC_PARAM_TC_ACC16) or 1 and fall back to 3 on the device when the refine of the
                                   uncertified rows would cost more.  Final labels are the same exact fp32 argmin
                                   whatever the tier */
#define KMC_PARAM_PROFILE   2   /* non-zero: record CUDA events around each phase (see below) */
#define KMC_PARAM_REFINE_KERNEL 4 /* exact refine of uncertified rows: 4 = warp per row (long candidate walks),
                                   8 = eight lanes per row (short walks), 0 (default) = both enqueued, the
                                   device picks from the walk lengths of the previous launch */
#define KMC_PARAM_CHUNKS    5   /* tensor path: row chunks per kmc_accumulate (1..4).  With more than one chunk the
                                   exact refine of chunk c runs on a side stream while the tensor kernel assigns
                                   chunk c+1; 1 (default) = strictly one kernel after the other */
#define KMC_PARAM_INCREMENTAL 6 /* 1 (default): kmc_accumulate updates running sums with the changed rows only where
                                   it can (see kmc_accumulate); 0: every pass sums every row */
#define KMC_PARAM_TC_PAIR   7   /* tensor path, 1-term kernel: 1 (default) = CTA pairs share every MMA (cta_group::2,
                                   M = 256: each SM stages half of every centroid tile); 0 = one CTA per MMA */
#define KMC_PARAM_TC_TILE_N 8   /* CTA-pair kernel: centroids per accumulator tile, 256 (default, two TMEM buffers) or 128
                                   (four buffers; measured slower at configs 4 and 5, kept for A/B runs) */
#define KMC_PARAM_DELTA_ATOMIC_ROWS 9 /* incremental accumulation: an iteration that changed at most this many labels
                                   applies the changed rows to the running sums with fp64 atomics (one launch) instead
                                   of the sort + segmented sum (ten); 0 = never, negative = the default (2^21 / D) */
#define KMC_PARAM_TC_ACC16 10  /* 1 (default): the adaptive policy's first tier uses fp16 accumulators where K > 256;
                                   0: fp32 accumulators everywhere */
int kmc_set_param(kmc_ctx* ctx, int param, double value);

/*
 * Device time per phase, averaged over the kmc_accumulate / kmc_update calls made since the last
 * call of this function (at most the 32 most recent), in milliseconds, measured with CUDA events
 * on the stream the kernels were launched on (needs KMC_PARAM_PROFILE).  Synchronises the events.
 *   out_ms[0] assignment kernel (fused small kernel / SIMT assign / tcgen05 assign)
 *   out_ms[1] exact refine of ambiguous rows (tensor path)      out_ms[2] sorted accumulation
 *   out_ms[3] centroid update (+ centroid plane preparation)    out_ms[4] calls averaged
 * Replaces the reference's external nvprof log parsing (scripts/compileResults.py:58-137) for
 * the phases of this path.
 */
#define KMC_PHASE_COUNT 5
int kmc_get_phase_ms(kmc_ctx* ctx, float* out_ms, int n);

/*
 * Debug / test hook of the tensor path: the raw approximate scores
 * s[i,k] = |c_k|^2 - 2 x_i . c_k (fp32, [n_rows, K] row-major) as the tcgen05 kernel sees them
 * before the argmin.  Only for small shards (n_rows * K * 4 bytes of caller memory).
 * No counterpart in the reference (which materialises the full distance matrix at :230).
 */
int kmc_debug_tc_scores(kmc_ctx* ctx, float* dev_scores, void* stream);

/*
 * Seeded k-means++ (D^2 sampling) on the device: K rows of dev_points [n_rows, D] are drawn into
 * dev_centroids [K, D]; the rows chosen are optionally returned in host_chosen_rows [K].  The same
 * seed gives the same centres (the K uniform numbers are a splitmix64 stream of `seed`).
 * dev_scratch needs kmc_kmeanspp_scratch_bytes(n_rows, K) bytes.  Needs no kmc_ctx.
 * Replaces `k_means_._init_centroids(data_batch, K, init='k-means++')` (reference :191), which is
 * host-side, private sklearn API and unseeded.
 */
int64_t kmc_kmeanspp_scratch_bytes(int64_t n_rows, int64_t K);
int kmc_kmeanspp(int device, const float* dev_points, int64_t n_rows, int64_t D, int64_t K, uint64_t seed,
                 float* dev_centroids, int64_t* host_chosen_rows, void* dev_scratch, int64_t scratch_bytes,
                 void* stream);

/*
 * Per-cluster sums, counts and inertia from labels the caller already has, WITHOUT any device
 * allocation and without a kmc_ctx or workspace.  No counterpart in the reference.
 *
 * For hosts that are not allowed to call cudaMalloc: an operator a framework invokes with nothing
 * but its input and output buffers (the framework's own pool owns all device memory and gives the
 * operator no scratch), or a latency-critical serving loop on a fixed memory budget.  The scratch
 * of the sorted accumulation (histogram, cursors, offsets, row permutation) lives in a fixed-size
 * __device__ array inside the library instead, about 4.3 MB, which every GPU that loads the
 * library holds until the process exits.  Limits: n_rows <= 1 048 576 and K <= 4096 per call;
 * larger sizes are rejected with KMC_E_NOMEM (use a kmc_ctx with a caller-provided workspace).
 *
 *   dev_points    fp32 [n_rows, D] row-major          dev_labels  int32 [n_rows], values in [0, K)
 *                                                      (rows with other labels are left out)
 *   dev_centroids fp32 [K, D] (for the inertia)       dev_partials double [K*D + K + 2], same layout
 *                                                      as kmc_partials_ptr() (sums | counts |
 *                                                      inertia | sum |x|^2); overwritten
 *
 * The first call on a device looks up the device address of that array (no allocation, but a
 * runtime call), so make the first call on each device outside any stream capture.  All calls on
 * one device share the array and it keeps the previous call's contents: the function clears what
 * must start at zero, and two calls whose kernels overlap on the same GPU (different streams or
 * threads) race on it -- serialise them.  Runs on `device` (an sm_100 GPU, checked on the first call) and puts
 * the calling thread's current device back before returning.
 */
int kmc_cluster_sums_noalloc(int device, const float* dev_points, const int32_t* dev_labels,
                             const float* dev_centroids, int64_t n_rows, int64_t D, int64_t K,
                             double* dev_partials, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KMC_B200_H_ */
