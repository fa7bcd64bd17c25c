/*
 * nnd_b200.h -- C ABI of libnnd_b200.so: brute-force 3-D nearest neighbour /
 * chamfer ("nnd") forward + backward for NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for the reference's native op layer
 * (dip/torch-nndistance/src/my_lib_cuda.cpp:25-41,44-72 and the launchers
 * dip/torch-nndistance/src/nnd_cuda.cu:132-162,185-222).  The reference binds
 * that layer through pybind11 with at::Tensor arguments; here the same calls
 * are plain pointers + sizes so any host language can bind them (ctypes stub in
 * INTEGRATION.md; the in-repo binding is pointcloudregistration_b200/_cabi.py).
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name says host;
 *   - clouds are row-major [b, n, 3] float32, contiguous (my_lib_cuda.cpp:5);
 *   - dist is squared L2 (no sqrt, my_lib.cpp:15), idx is int32 (torch_nndistance/__init__.py:19-20);
 *   - the caller owns every buffer, the library writes in place (nnd_cuda.cu:140-149);
 *   - `stream` is a cudaStream_t (NULL = legacy default stream, which is what the
 *     reference always uses, nnd_cuda.cu:152-153);
 *   - return value: 0 on success, otherwise a cudaError_t value or one of the
 *     NND_E* codes below.  (The reference returns 1 on success / 0 on failure,
 *     nnd_cuda.cu:155-161; the torch_nndistance_aten shim maps back to that.)
 *   - results: idx bit-exact and dist bit-exact vs the reference CPU path
 *     my_lib.cpp:3-25 (lowest index on ties, no FMA contraction); gradients
 *     within 1e-5 relative of my_lib.cpp:91-130 (atomic summation order).
 *   - there is NO CPU fallback in this library.
 *   - threads and streams: every call only enqueues work on `stream` and returns; calls may come from any host
 *     thread.  Forwards over >= 200 000 rows (the panel scan) fork into library-owned streams and join back into
 *     `stream` before returning, and share one constant-memory window per device: such forwards are serialised on
 *     the device even when they come from different streams.  A panel scan captured into a CUDA graph is ordered
 *     only by the stream that replays the graph; do not replay it concurrently with other large forwards.
 */
#ifndef NND_B200_H
#define NND_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NND_EINVAL      10001  /* bad argument (negative size, null pointer, ...)   */
#define NND_EWORKSPACE  10002  /* workspace too small or misaligned                 */
#define NND_EDEVICE     10003  /* device is not compute capability 10.x             */
#define NND_ETOOBIG     10004  /* beyond the built-in scratch of the no-allocation entry point */

/* Library / build identification, e.g. "nnd_b200 0.1.0 sm_100a". */
const char *nnd_version(void);

/* Human-readable text for a return code of this library. */
const char *nnd_error_string(int code);

/* Scratch bytes needed by nnd_forward_cuda_f32 for a [b,n,3] x [b,m,3] problem
 * (structure-of-arrays copies of both clouds + packed (dist,idx) keys).
 * The workspace must be 256-byte aligned device memory. */
size_t nnd_workspace_bytes(int b, int n, int m);

/* ---- row-range forward: the unit the multi-GPU path calls (new design; the
 * reference has no multi-device path).  One evaluation of every (xyz1[j], xyz2[k]) distance serves both
 * directions, so a rank that owns rows [row_lo, row_hi) of xyz1 produces
 *   - dist1/idx1 for its rows (final), written into the FULL-SIZE [b,n] buffers at those rows only, and
 *   - colkeys[b * nnd_colkeys_count(1, m)]: for every point of xyz2 the packed partial result
 *     (float_bits(dist) << 32 | row_group) over this rank's rows.  Keys of different ranks merge with an
 *     element-wise integer MIN (every key is < 2^63: signed and unsigned order agree), e.g. an NCCL
 *     all-reduce(MIN) on int64; nnd_finalize_cols_cuda_f32 then turns merged keys into dist2/idx2.
 * row_lo must be a multiple of nnd_row_group(); row_hi must be one too unless row_hi == n.
 * nnd_forward_cuda_f32 is exactly rows [0, n) followed by the column finalize. */
size_t nnd_rows_workspace_bytes(int b, int n, int m);
size_t nnd_colkeys_count(int b, int m);
int nnd_row_group(void);
int nnd_forward_rows_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                              int row_lo, int row_hi, float *dist1, int32_t *idx1,
                              unsigned long long *colkeys,
                              void *workspace, size_t workspace_bytes, void *stream);
int nnd_finalize_cols_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                               const unsigned long long *colkeys,
                               float *dist2, int32_t *idx2, void *stream);

/* ---- the multi-GPU exchange over PEER MEMORY (new design; the reference has no multi-device path).  One process per GPU;
 * every rank holds its column keys, dist1 / idx1 (own rows final) and output buffers in memory the other ranks can address
 * (NVLink / NVSwitch P2P: e.g. torch.distributed._symmetric_memory), and passes HOST arrays of the `world` device pointers,
 * its own included, in rank order.  nnd_exchange_cols_p2p_cuda_f32 is ONE kernel for "all_reduce(MIN) of the column keys +
 * nnd_finalize_cols_cuda_f32 + all_gather of dist1 / idx1": rank r merges and finalizes its slice of the columns reading
 * every rank's keys directly, stores dist2 / idx2 of that slice into every rank's buffers and pushes its own rows
 * [row_lo, row_hi) of dist1 / idx1 to every rank.  nnd_sum_p2p_cuda_f32 sums `count` floats of every rank's buffer in
 * place (two-shot, rank order: the same bits on every rank and run to run).  THE CALLER brackets each launch with a
 * barrier over all ranks on `stream` (before: every rank's inputs are complete; after: every store has landed).
 * At most nnd_peer_max_ranks() ranks.  Results are bit-identical to the NCCL formulation for any world size. */
int nnd_peer_max_ranks(void);
int nnd_exchange_cols_p2p_cuda_f32(const void *const *peer_colkeys, void *const *peer_dist1, void *const *peer_idx1,
                                   void *const *peer_dist2, void *const *peer_idx2, int world, int rank,
                                   const float *xyz1, const float *xyz2, int b, int n, int m,
                                   int row_lo, int row_hi, void *stream);
int nnd_sum_p2p_cuda_f32(void *const *peer_bufs, int world, int rank, long long count, void *stream);

/* Scratch bytes needed by nnd_query_cuda_f32 (one direction). */
size_t nnd_query_workspace_bytes(int b, int nq, int nt);

/* Replaces nnd_forward_cuda (my_lib_cuda.cpp:25-41) -> NmDistanceKernelLauncher
 * (nnd_cuda.cu:132-162): both directions.
 *   dist1[b,n], idx1[b,n] : for every point of xyz1 its nearest point of xyz2
 *   dist2[b,m], idx2[b,m] : for every point of xyz2 its nearest point of xyz1
 * m == 0 (resp. n == 0) gives dist = 0, idx = 0 like my_lib.cpp:9-10,21-22. */
int nnd_forward_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                         float *dist1, float *dist2, int32_t *idx1, int32_t *idx2,
                         void *workspace, size_t workspace_bytes, void *stream);

/* Which forward nnd_forward_cuda_f32 runs for a batch of b cloud pairs of n and m points:
 *   0 = single-launch small-cloud kernel (no workspace; 2*b*n*m evaluations): both clouds <= nnd_small_max_points(),
 *       or -- a lone pair, where the general path cannot fill the machine -- up to 2.5x that with b*max(n,m) <= 3x that,
 *   1 = two-pass kernel (prepare + scan + finalize; 2*b*n*m evaluations),
 *   2 = symmetric kernels (every distance evaluated once: b*n*m evaluations).  Results are bit-identical on all. */
int nnd_forward_kind(int b, int n, int m);

/* Small clouds (the sizes the reference's callers really use: DIP patches B=256 x 256 points, dip/train.py:84; NDP
 * clouds of ~3 k points, registration.py:236).  When both clouds have at most nnd_small_max_points() points the forward
 * is ONE launch (whole target cloud in shared memory, index resolved in the same kernel, no workspace: pass NULL / 0)
 * and the backward is ONE launch (every CTA owns a range of gradient rows in shared memory: no memsets, no global
 * atomics).  nnd_workspace_bytes and
 * nnd_query_workspace_bytes return 0 for such shapes.  nnd_set_small_max_points(points >= 0) sets the limit and returns
 * the previous one (negative: query only); default 4096, environment NND_B200_SMALL_MAX, hard maximum 16384. */
int nnd_small_max_points(void);
int nnd_set_small_max_points(int points);
/* Tuning / test knob of the single-launch forward: warps per CTA (2, 4, 8, 16), queries per lane (1, 2, 4, 8) and how many
 * warps share one set of queries while splitting the target cloud between them (a divisor of warps).  (0, 0, 0)
 * restores the default (chosen from the problem size).  Results are bit-identical for every setting. */
int nnd_set_small_config(int warps, int qpt, int tsplit);

/* Ragged batches: per-cloud valid lengths len1[b], len2[b] (int64 DEVICE arrays, NULL = all points), the
 * `lengths1/lengths2` of pytorch3d's knn_points as c2p-net/deformationpyramid/model/loss.py:143-144 calls it.
 * Neighbours are searched among the first len points of the other cloud only; rows at or past a cloud's length get
 * dist = 0, idx = 0 (gradient 0); a cloud of length 0 gives dist = 0, idx = 0 for its partner (my_lib.cpp:9-10).
 * Lengths are clamped to [0, n] by the kernel.  One launch each, whole batch.  Only for shapes with
 * nnd_ragged_supported(n, m) == 1 (single-launch kernels: both clouds <= nnd_small_max_points()); otherwise
 * NND_EINVAL and the caller loops over the batch.  nnd_backward_ragged / nnd_knn_backward_ragged OVERWRITE both
 * gradients (zero init fused). */
int nnd_ragged_supported(int n, int m);
int nnd_forward_ragged_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                                const long long *len1, const long long *len2,
                                float *dist1, float *dist2, int32_t *idx1, int32_t *idx2, void *stream);
int nnd_query_ragged_cuda_f32(const float *query, const float *target, int b, int nq, int nt,
                              const long long *lenq, const long long *lent,
                              float *dist, int32_t *idx, void *stream);
int nnd_backward_ragged_cuda_f32(const float *xyz1, const float *xyz2,
                                 const float *graddist1, const float *graddist2,
                                 const int32_t *idx1, const int32_t *idx2, int b, int n, int m,
                                 const long long *len1, const long long *len2,
                                 float *gradxyz1, float *gradxyz2, void *stream);
int nnd_knn_backward_ragged_cuda_f32(const float *query, const float *target, const float *graddist,
                                     const int32_t *idx, int b, int nq, int nt,
                                     const long long *lenq, const long long *lent,
                                     float *grad_query, float *grad_target, void *stream);

/* Warm start for registration loops (new; the reference recomputes from scratch every iteration,
 * c2p-net/deformationpyramid/model/registration.py:231-261).  Same call as nnd_forward_cuda_f32 plus
 * hint_idx2[b, m]: for every point of xyz2 a row of xyz1 that was its neighbour in the PREVIOUS iteration (typically
 * the idx2 the last call returned; it MAY be the same buffer as idx2).  The symmetric kernels use the distance to that
 * row, evaluated now, as the initial upper bound of the column minimum, which removes almost all of their column
 * bookkeeping when the clouds moved little.  RESULTS DO NOT DEPEND ON THE HINT (any values are safe: out-of-range
 * indices are ignored); NULL = nnd_forward_cuda_f32.  Small clouds and the two-pass path ignore it. */
int nnd_forward_warm_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                              float *dist1, float *dist2, int32_t *idx1, int32_t *idx2,
                              const int32_t *hint_idx2,
                              void *workspace, size_t workspace_bytes, void *stream);

/* EXACT search with spatial pruning (opt-in; new design -- the reference only has the brute-force scan my_lib.cpp:3-25 /
 * nnd_cuda.cu:8-130, which nnd_forward_cuda_f32 replaces and on which this project's metric is defined).  Same arguments
 * and THE SAME BITS as nnd_forward_cuda_f32 on every input (dist bit-exact, lowest index on ties, the NaN / inf rules of
 * my_lib.cpp:16), but only the pairs that can matter are evaluated: both clouds are sorted along a Morton curve, cut
 * into tiles with bounding boxes, and a warp of 32 neighbouring queries skips every tile whose distance lower bound
 * (shrunk to cover fp32 rounding) exceeds its current best (csrc/nnd_pruned.cuh).  The time depends on the geometry:
 * clouds that overlap (registration inputs) run one to two orders of magnitude faster than the brute-force scan at
 * 10^5 .. 10^6 points; degenerate inputs where every distance ties (all points identical) prune nothing and run
 * slower than nnd_forward_cuda_f32.  `workspace` = nnd_pruned_workspace_bytes(b, n, m) of 256-byte aligned device
 * memory (0 = b*n or b*m exceeds 2^31 - 1; also b <= 65535).  Uses cub::DeviceRadixSort for the curve order. */
size_t nnd_pruned_workspace_bytes(int b, int n, int m);
int nnd_forward_pruned_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                                float *dist1, float *dist2, int32_t *idx1, int32_t *idx2,
                                void *workspace, size_t workspace_bytes, void *stream);
/* The pruned search inside a registration loop (registration.py:231-261: one cloud moves a little per iteration, the other
 * never): nnd_forward_pruned_cuda_f32 with parts of the PREVIOUS call's preparation reused.  `reuse` bits:
 *   1 = keep xyz1's curve ORDER from the previous call (skip its sort).  Its points and every box are rebuilt from the
 *       CURRENT coordinates, so the results stay exact however far the cloud has moved -- a stale order only costs pruning
 *       power; re-sort (clear the bit) every few dozen iterations;
 *   2 = xyz2 is bit-for-bit unchanged since the previous call: all of its preparation is reused as it is.
 * The workspace must be the one the previous call used, untouched in between, with the same b, n, m (n, m > 0).
 * reuse = 0 is nnd_forward_pruned_cuda_f32.  RESULTS ARE THE SAME BITS for every value of `reuse` that is used as described. */
int nnd_forward_pruned_reuse_cuda_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                                      float *dist1, float *dist2, int32_t *idx1, int32_t *idx2,
                                      void *workspace, size_t workspace_bytes, int reuse, void *stream);
/* Tuning / test knob of the pruned search: whether the 4 warps of a CTA share one group of 32 queries (each taking a
 * quarter of the target tiles) or own a group each.  0 = chosen from the problem size (default: shared below 16 384
 * groups per call and for target clouds of at least 65 536 points), 1 = never, 2 = always.  Results are bit-identical.
 * Returns the previous mode; other values only query. */
int nnd_set_pruned_share(int mode);
/* One direction of the same search (the pruned twin of nnd_query_cuda_f32): for every query point its nearest target
 * point.  workspace = nnd_pruned_workspace_bytes(b, nq, nt). */
int nnd_query_pruned_cuda_f32(const float *query, const float *target, int b, int nq, int nt,
                              float *dist, int32_t *idx,
                              void *workspace, size_t workspace_bytes, void *stream);

/* 1 when nnd_forward_cuda_f32 takes the symmetric kernel for clouds of n and m points (every distance evaluated
 * once: b*n*m evaluations per call), 0 when it takes the two-pass kernel (2*b*n*m evaluations; small clouds).
 * Results are bit-identical either way; bench.py uses this to count the work a launch executes. */
int nnd_forward_is_symmetric(int n, int m);

/* Measurement aid for bench.py's roofline block.  While enabled (per calling thread), every forward /
 * query call also records CUDA events on its stream around each of its three kernels (no
 * synchronisation, up to 64 calls are remembered).  nnd_profile_collect waits for those events and
 * returns the summed kernel durations in milliseconds since the last collect:
 * kernel_ms[0] = prepare (AoS->SoA), [1] = scan (the dominant kernel), [2] = finalize (key unpack),
 * and the number of calls they cover. */
int nnd_profile_enable(int on);
int nnd_profile_collect(float *kernel_ms, int *calls);

/* One direction only (the unit the query-sharded multi-GPU path calls):
 * for every query point its nearest target point. */
int nnd_query_cuda_f32(const float *query, const float *target, int b, int nq, int nt,
                       float *dist, int32_t *idx,
                       void *workspace, size_t workspace_bytes, void *stream);

/* No-allocation variant of nnd_query_cuda_f32.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and bit-exact results as nnd_query_cuda_f32, but the scratch (structure-of-
 * arrays copies of both clouds + packed keys) lives in a static device array inside the library, and the call makes no
 * cudaMalloc / cudaFree / allocating library call.
 *   - limit: nnd_query_workspace_bytes(b, nq, nt) <= NND_NOALLOC_SCRATCH_BYTES (8 MiB, = nnd_query_noalloc_max_bytes();
 *     12 B per padded-to-512 point of both clouds + 8 B per query, e.g. b * nq = b * nt = 262 144 fits exactly).
 *     Larger problems return NND_ETOOBIG and launch nothing.  Shapes of the single-launch kernels (workspace 0) are
 *     accepted too and do not touch the array.
 *   - every device that loads the library holds the 8 MiB until the process exits;
 *   - every call on a device shares the one array: calls that OVERLAP on the GPU (different streams, or a graph
 *     replay beside an eager call) RACE -- keep them in one stream's order.  Nothing in the array has to be cleared
 *     between calls: each call rewrites what it reads (the previous call's contents are harmless). */
#define NND_NOALLOC_SCRATCH_BYTES (8u << 20)
size_t nnd_query_noalloc_max_bytes(void);
int nnd_query_noalloc_cuda_f32(const float *query, const float *target, int b, int nq, int nt,
                               float *dist, int32_t *idx, void *stream);

/* Replaces nnd_backward_cuda (my_lib_cuda.cpp:44-72) -> NmDistanceGradKernelLauncher
 * (nnd_cuda.cu:185-222).  OVERWRITES gradxyz1[b,n,3] and gradxyz2[b,m,3] (zero
 * init fused, like the CPU path my_lib.cpp:91-94; the reference CUDA path
 * accumulates into caller-zeroed buffers, nnd_cuda.cu:199-200 -- identical when
 * the caller passes zeros, which torch_nndistance/__init__.py:42-43 does). */
int nnd_backward_cuda_f32(const float *xyz1, const float *xyz2,
                          const float *graddist1, const float *graddist2,
                          const int32_t *idx1, const int32_t *idx2,
                          int b, int n, int m,
                          float *gradxyz1, float *gradxyz2, void *stream);

/* Deterministic backward (opt-in; the reference has no such mode on the GPU -- its atomicAdd order is arbitrary,
 * nnd_cuda.cu:175-180).  Same arguments and the same OVERWRITE semantics as nnd_backward_cuda_f32, but the gradients are
 * BIT-IDENTICAL to the reference CPU path my_lib.cpp:91-130 (its sequential accumulation order is reproduced: sort of
 * (neighbour row, own row) keys + one in-order pass per gradient row; no atomics, run-to-run deterministic).
 * `workspace` = nnd_backward_exact_workspace_bytes(b, n, m) of 256-byte aligned device memory (0 = b*n or b*m
 * exceeds 2^31 - 1, which this entry point rejects with NND_EINVAL).  Slower than the atomic kernel (measured at
 * N = M = 1M: 0.41 ms against 0.06 ms on a uniform cloud, 2.5 ms against 0.15 ms on a clustered one, where thousands of
 * rows share a neighbour and each such list has to be summed in order; profiles/r01_bwd_exact.txt) -- still
 * 0.2 % / 1 % of the forward. */
size_t nnd_backward_exact_workspace_bytes(int b, int n, int m);
int nnd_backward_exact_cuda_f32(const float *xyz1, const float *xyz2,
                                const float *graddist1, const float *graddist2,
                                const int32_t *idx1, const int32_t *idx2,
                                int b, int n, int m,
                                float *gradxyz1, float *gradxyz2,
                                void *workspace, size_t workspace_bytes, void *stream);

/* Tuning knob.  The symmetric forward has two scan kernels: the persistent TMA-fed one and, for row ranges of at least
 * this many rows, the "panel" one that keeps the column points in uniform registers (one launch per panel of columns,
 * DESIGN.md section 4.0).  Results are bit-identical.  Sets the threshold (rows > 0) and returns the previous value;
 * rows <= 0 only queries.  Default 200000, or the environment variable NND_B200_PANEL_MIN_ROWS. */
long long nnd_set_panel_min_rows(long long rows);
/* Which scan nnd_forward_rows_cuda_f32 runs for rows [row_lo, row_hi) of a [b, n]-point xyz1 against a [b, m]-point
 * xyz2: 1 = persistent symmetric kernel, 2 = panel kernel, 3 = panel kernel with the roles of the clouds swapped
 * (ranges shorter than m), 0 = nothing to do.  (nnd_forward_cuda_f32 = the full range when
 * nnd_forward_is_symmetric(n, m).)  The same helper decides inside nnd_forward_rows_cuda_f32.  One exception: a call
 * whose stream is being captured into a CUDA graph always takes the persistent kernel (results are bit-identical). */
int nnd_rows_scan_kind(int b, int n, int m, int row_lo, int row_hi);
/* Automatic hints for the column filter (cold start; csrc/nnd_hint.cuh): a hash grid of the register-side cloud gives
 * every streamed point a nearby row, whose exact distance seeds the filter -- the same mechanism as the caller-supplied
 * hint of nnd_forward_warm_cuda_f32, which takes precedence.  0 = off, 1 = in front of the panel scans (default), 2 = in
 * front of the persistent symmetric kernel as well.  Results do not depend on it.  Returns the previous mode; other
 * values only query. */
int nnd_set_auto_hint(int mode);
/* Tuning knob: sample tiles (512 rows each) of the column-filter seed pass run in front of the PERSISTENT symmetric kernel
 * (default 1; 0 = off; the panel scan always seeds).  Results do not depend on it.  Returns the previous value; negative
 * only queries. */
int nnd_set_sym_seed_tiles(int tiles);
/* Column tiles (512 points) per panel launch: 0 = chosen from the row count (default), 1, 2 or 4 = forced.  Returns the
 * previous setting; other values only query.  Environment: NND_B200_PANEL_TILES. */
int nnd_set_panel_tiles(int tiles);

/* Fused truncated-chamfer epilogue (c2p-net/deformationpyramid/model/loss.py:150-215 with mean/mean reductions, the
 * form registration.py:236 uses):  loss[0] = scale1 * sum(dist1 < trunc ? dist1 : 0) + scale2 * sum(dist2 < trunc ? dist2 : 0),
 * and, when the pointers are non-null, graddist1[i] = dist1[i] < trunc ? scale1 : 0 (likewise graddist2): d loss / d dist,
 * ready for nnd_backward_cuda_f32.  Deterministic (fixed summation order).  `workspace` is
 * nnd_chamfer_reduce_workspace_bytes() of 256-byte aligned device memory; it needs no initialisation and carries no
 * state between calls (any scratch will do, e.g. a fresh allocation per call). */
size_t nnd_chamfer_reduce_workspace_bytes(void);
int nnd_chamfer_reduce_cuda_f32(const float *dist1, const float *dist2, int b, int n, int m,
                                float trunc, float scale1, float scale2,
                                float *loss, float *graddist1, float *graddist2,
                                void *workspace, size_t workspace_bytes, void *stream);

/* One term of the backward (the unit the query-sharded multi-GPU path calls):
 * for query rows [0,nq) with nearest targets idx:  t = 2*graddist*(q - target[idx]);
 * grad_query[row] += t ; grad_target[idx] -= t.  ACCUMULATES (caller zeroes). */
int nnd_backward_term_cuda_f32(const float *query, const float *target,
                               const float *graddist, const int32_t *idx,
                               int b, int nq, int nt,
                               float *grad_query, float *grad_target, void *stream);

/* Number of kernels the last forward / backward call on this thread launched
 * (bench.py's gpu_launches accounting). */
int nnd_last_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* NND_B200_H */
