/* surfel_b200 -- C ABI of the B200-native 2D-Gaussian (surfel) differentiable rasterizer.
 *
 * Drop-in boundary for VINGS-Mono's rasterizer extension.  These entry points are what the
 * reference's torch glue (submodules/diff-surfel-rasterization/rasterize_points.cu) would call
 * instead of CudaRasterizer::Rasterizer::{forward,backward,markVisible} and ADAM::adamUpdate:
 *
 *   sb200_forward_begin + sb200_forward_finish  <->  RasterizeGaussiansCUDA          rasterize_points.cu:41-149
 *                                                    Rasterizer::forward             cuda_rasterizer/rasterizer_impl.cu:305-478
 *   sb200_backward                              <->  RasterizeGaussiansBackwardCUDA  rasterize_points.cu:151-256
 *                                                    Rasterizer::backward            cuda_rasterizer/rasterizer_impl.cu:482-599
 *   sb200_mark_visible                          <->  markVisible                     rasterize_points.cu:258-277
 *   sb200_adam_update                           <->  adamUpdate                      rasterize_points.cu:281-306, cuda_rasterizer/adam.cu:40-67
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous fp32/int32/u8 data unless it says "host";
 *   - the library never allocates or frees device memory: workspaces are caller-allocated from the
 *     sizes returned by the *_bytes queries (the reference grows torch tensors through
 *     std::function<char*(size_t)> callbacks, rasterizer.h:32-35; here the caller does the resize
 *     between forward_begin and forward_finish, which is the only point a size is data dependent);
 *   - every launch goes to the explicit `stream` (the reference uses the legacy default stream);
 *   - the only host synchronisation is inside sb200_forward_begin (one 8-byte D2H read of the pair
 *     count R; the reference blocks twice, rasterizer_impl.cu:391,439);
 *   - return value 0 = ok, non-zero = error, message via sb200_last_error(); nothing is thrown.
 *   - matrices are the reference's: 16 floats, the row-major torch tensor holding the transposed
 *     (column-major-read) world->view and full projection matrices (auxiliary.h:91-130).
 */
#ifndef SURFEL_B200_H
#define SURFEL_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ABI version of this header (bumped on any signature change). */
int sb200_abi_version(void);
/* Message of the last failing call on this thread ("" if none). */
const char* sb200_last_error(void);

/* ---- measurement hooks (bench.py) ------------------------------------------------------------
 * Number of kernels this library has launched since load (the `gpu_launches` claim of bench.py). */
long long sb200_launch_count(void);
/* Per-stage device timing: while enabled, every stage is bracketed by a cudaEvent pair recorded on the
 * stream the stage is launched on.  sb200_profile_read waits for the recorded events, returns the summed
 * milliseconds and the number of timed launches per stage since the previous read, and resets. */
enum {
    SB200_STAGE_PREPROCESS = 0,     /* per-surfel preprocess (+ R) */
    SB200_STAGE_DUPLICATE = 1,      /* (tile, surfel) pair emission in depth order + pair-offset scan + tile counts */
    SB200_STAGE_SORT = 2,           /* stable onesweep sort of the pairs by tile id */
    SB200_STAGE_RANGES = 3,         /* tile ranges, tile launch order, digit bases of the pair sort */
    SB200_STAGE_BLEND_FWD = 4,      /* per-tile forward blend */
    SB200_STAGE_BLEND_BWD = 5,      /* per-tile backward blend (+ zeroing of its accumulation record) */
    SB200_STAGE_PREPROCESS_BWD = 6, /* per-surfel backward preprocess */
    SB200_STAGE_ADAM = 7,           /* masked Adam update */
    SB200_STAGE_MAP_LOSS = 8,       /* fused mapping loss (5 kernels) */
    SB200_STAGE_ACTIVATE = 9,       /* raw parameters -> rasterizer inputs */
    SB200_STAGE_UPDATE = 10,        /* fused per-surfel update (activation backward, records, weighting, Adam) */
    SB200_STAGE_CHURN = 11,         /* surfel add / prune: kNN distances, stable compaction */
    SB200_STAGE_SURFEL_SORT = 12,   /* stable onesweep sort of the visible surfels by depth (histogram + 4 passes) */
    SB200_STAGE_DP_REDUCE = 13,     /* multi-GPU: reduce-scatter of the stage rows by peer loads (sb200_dp_reduce) */
    SB200_STAGE_DP_POST = 14,       /* multi-GPU: all-gather of the result rows fused into bookkeeping + Adam (sb200_dp_post) */
    SB200_NUM_STAGES = 15
};
int sb200_profile_enable(int on);
int sb200_profile_read(double ms[SB200_NUM_STAGES], long long count[SB200_NUM_STAGES]);

/* ---- workspace sizes ------------------------------------------------------------------------ */
/* per-surfel workspace (reference: GeometryState, rasterizer_impl.h:29-44); must persist forward->backward */
size_t sb200_geom_bytes(int64_t P);
/* per-image workspace: tile ranges + per-pixel final T/M1/M2 and contributor counts
 * (reference: ImageState, rasterizer_impl.h:46-65); must persist forward->backward */
size_t sb200_image_bytes(int W, int H);
/* sort scratch for R (tile,surfel) pairs (reference: BinningState, rasterizer_impl.h:86-97); dead after forward_finish */
size_t sb200_binning_bytes(int64_t R);
/* the pair list: u32 sorted surfel ids [R] followed by the forward's hit masks (one u32 per 32-surfel batch of a
 * tile and warp of its CTA: which surfels of the batch some live pixel of the warp's 8x4 block may blend; the
 * backward visits exactly those).  Written by sb200_forward_finish / _async, read by sb200_backward with the SAME
 * R (or pair_capacity); must persist forward->backward.  Replaces the reference's per-bucket checkpoints
 * (SampleState, rasterizer_impl.h:68-83: 13 KiB per 32 surfels) with ~1 byte per pair. */
size_t sb200_pair_list_bytes(int64_t R, int W, int H);
/* per-surfel gradient accumulation scratch for the backward: 80 B * P, zeroed by the library */
size_t sb200_grad_bytes(int64_t P);
/* byte offsets of the inspectable arrays inside the workspaces (tests / debugging):
 *   geom:    [0] records float4[8P]  [1] depth f32[P]  [2] tile rect u16x4[P]  [3] tiles_touched u32[P]
 *            [4] ids of the V visible surfels in (depth, id) order u32[P] (entries past V are scratch)
 *   image:   [0] ranges uint2[tiles] [1] final T,M1,M2 f32[3HW]  [2] n_contrib,median_contrib u32[2HW]
 *   binning: [0] tile ids A u32[R] (pairs as emitted, depth order)  [1] tile ids B u32[R]
 *            [2] surfel ids A u32[R] (as emitted)  [3] surfel ids B u32[R] */
void sb200_geom_layout(int64_t P, size_t out[5]);
void sb200_image_layout(int W, int H, size_t out[3]);
void sb200_binning_layout(int64_t R, size_t out[4]);

/* ---- forward -------------------------------------------------------------------------------- */
/* Stage 1: per-surfel preprocess (ray-splat transform, AABB, radius, tile rect, R = sum of the tile counts) and
 * the stable sort of the visible surfels by depth.  Writes radii[P] (int32) and the geometry workspace.  Waits
 * ONCE, for the 8-byte read-back of R into *num_rendered_host (host pointer, pinned preferred), which is queued
 * before the sort so that the GPU keeps working while the host sizes the binning scratch. */
int sb200_forward_begin(int P, int W, int H,
                        const float* means3D,    /* [P,3] */
                        const float* scales,     /* [P,2] */
                        const float* rotations,  /* [P,4] (w,x,y,z) */
                        const float* opacities,  /* [P]   */
                        const float* colors,     /* [P,3] colors_precomp */
                        const float* viewmatrix, const float* projmatrix,
                        void* geom_ws, int* radii, int64_t* num_rendered_host, void* stream);

/* Stage 2: pair emission in depth order, tile ranges, stable sort of the pairs by tile, per-tile blend.
 * point_list (sb200_pair_list_bytes(R, W, H) bytes) receives the sorted surfel ids [R] and, behind them, the hit
 * masks; it must persist until the backward.
 * out_color [3,H,W]; out_others [7,H,W] = depth, alpha, normal xyz (view space), median depth, distortion. */
int sb200_forward_finish(int P, int W, int H, int64_t R, const float* bg /* [3] */,
                         void* geom_ws, void* binning_ws, uint32_t* point_list, void* image_ws,
                         float* out_color, float* out_others, void* stream);

/* Both stages in one call WITHOUT a host synchronisation, for callers that keep a binning scratch of
 * `pair_capacity` pairs (sb200_binning_bytes(pair_capacity), sb200_pair_list_bytes(pair_capacity, W, H)): every launch takes the
 * pair count from device memory.  R is copied to *num_rendered_host (pinned host memory, owned by the caller's
 * workspace) right after the preprocess and `sync` (a handle from sb200_sync_create, one per workspace) is recorded
 * behind that copy; sb200_forward_wait(sync) blocks until the copy has landed - call it wherever convenient before
 * committing the view's results (the mapper does just before its parameter update, with the rest of the view
 * already queued, so the GPU never idles).  Count buffer and handle belong to ONE workspace: forwards of other
 * workspaces, streams or threads in between do not disturb them (re-entrant per workspace).
 * If R > pair_capacity only the first pair_capacity pairs (depth order) were binned: the outputs are memory-safe
 * but wrong, and the caller must grow the scratch and run the forward again.  Pass pair_capacity (not R) as R to
 * sb200_backward. */
int sb200_forward_async(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                        const float* opacities, const float* colors, const float* viewmatrix, const float* projmatrix,
                        const float* bg, void* geom_ws, int* radii, void* binning_ws, int64_t pair_capacity,
                        uint32_t* point_list, void* image_ws, float* out_color, float* out_others,
                        int64_t* num_rendered_host, void* sync, void* stream);
int sb200_forward_wait(void* sync);
/* hand-off handles of sb200_forward_async (a CUDA event of the current device); NULL on failure */
void* sb200_sync_create(void);
void sb200_sync_destroy(void* sync);

/* ---- backward ------------------------------------------------------------------------------- */
/* All dL_d* outputs and scores are fully written (no pre-zeroing needed).  dL_dtransMat may be NULL.
 * pixel_mask: u8/bool [H*W]; masked-out pixels contribute no gradient and no score. */
int sb200_backward(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                   const float* means3D, const float* scales, const float* rotations,
                   const float* viewmatrix, const float* projmatrix, const int* radii,
                   const void* geom_ws, const uint32_t* point_list, const void* image_ws,
                   const float* out_color, const float* out_others,
                   const float* dL_dout_color /* [3,H,W] */, const float* dL_dout_others /* [7,H,W] */,
                   const uint8_t* pixel_mask, void* grad_ws,
                   float* dL_dmeans3D /* [P,3] */, float* dL_dmeans2D /* [P,3] */, float* dL_dcolors /* [P,3] */,
                   float* dL_dopacity /* [P,1] */, float* dL_dscales /* [P,2] */, float* dL_drotations /* [P,4] */,
                   float* dL_dtransMat /* [P,9] or NULL */, float* scores /* [P,2] */, void* stream);

/* ---- auxiliaries exported by the same reference module -------------------------------------- */
int sb200_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present /* bool [P] */, void* stream);
int sb200_adam_update(float* param, const float* param_grad, float* exp_avg, float* exp_avg_sq,
                      const uint8_t* visible /* bool [N] */, float lr, float b1, float b2, float eps,
                      uint32_t N, uint32_t M, void* stream);

/* ---- the mapping step either side of the rasterizer -----------
 * These replace torch code of the CALLER (scripts/gaussian/), not entry points of the reference extension;
 * the reference-side binding is the Python in vings-mono_b200/diff_surfel_rasterization/mapping.py.
 *
 * sb200_map_loss: get_loss (scripts/gaussian/loss_utils.py:110-153, use_sky False) on the rasterizer's raw
 * outputs, including render_raw's post-processing (scripts/gaussian/gaussian_base.py:205-221: depth =
 * nan_to_num(allmap0/allmap1), normals rotated by w2c[:3,:3], depth_propagate_normal of
 * scripts/gaussian/normal_utils.py:6-21) and the backward of all of it.
 *   color [3,H,W], allmap [7,H,W]: sb200_forward_finish outputs; gt_rgb [3,H,W], gt_depth [H,W], gt_depth_cov [H,W];
 *   w2c: 16 floats, row-major 4x4 world->camera pose (device);
 *   (cx, cy, fx, fy): camera point of pixel (col, row) at depth d is ((col-cx)/fx*d, (row-cy)/fy*d, d).  In the
 *   reference's naming u is the ROW: cx = intrinsic cv, fx = fv, cy = cu, fy = fu (scripts/gaussian/tf.py:139-143,
 *   configs/rtg/hotel.yaml:31-34);
 *   ws: sb200_map_loss_bytes(W, H) bytes of scratch;
 *   loss [7] (device): total, then the six means (L1, SSIM, normal, alpha, depth, dist; normal is 0 when
 *   w_normal == 0).  Means over an empty pixel set are NaN, like torch's;
 *   dL_dcolor [3,H,W], dL_dallmap [7,H,W]: fully written.  Where allmap1 == 0 torch's autograd yields NaN for
 *   dL_dallmap[0:2] (0/0); no surfel is blended there, and 0 is written instead. */
size_t sb200_map_loss_bytes(int W, int H);
int sb200_map_loss(int W, int H, const float* color, const float* allmap, const float* gt_rgb, const float* gt_depth,
                   const float* gt_depth_cov, const float* w2c, float cx, float cy, float fx, float fy,
                   float w_rgb, float w_normal, float w_alpha, float w_depth, float w_dist,
                   void* ws, float* loss, float* dL_dcolor, float* dL_dallmap, void* stream);

/* sb200_activate: opacities = sigmoid(opacity_raw) [P], scales = exp(scaling_raw) [P,2], rotations =
 * normalize(rotation_raw) [P,4] (scripts/gaussian/gaussian_base.py:160-166 via the model's activations). */
int sb200_activate(int P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                   float* opacities, float* scales, float* rotations, void* stream);

/* sb200_surfel_update: one pass over the surfels after sb200_backward: backward of the three activations,
 * add_records (scripts/gaussian/gaussian_model.py:163-168, gaussian_base.py:388-392), weighting_grad
 * (scripts/gaussian/gaussian_utils.py:56-62), radii[_stable_mask] = 0 and SparseGaussianAdam.step(radii > 0)
 * (gaussian_base.py:400-403) for the five groups, in the order xyz[3], rgb[3], opacity[1], scaling[2], rotation[4].
 *   params / exp_avg / exp_avg_sq: HOST arrays of 5 device pointers; lr: HOST array of 5 floats;
 *   d_*: sb200_backward outputs (gradients w.r.t. the activated inputs); cur_scores [P,2] its scores output;
 *   local_scores, global_scores [P,2], kf_max_scores [P], kf_id int64 [P]: the mapper's score bookkeeping; kf =
 *   id of the current keyframe.
 * mode SB200_UPDATE_LOCAL does all of it.  For views sharded over ranks, SB200_UPDATE_DP_PRE writes the weighted
 * raw-parameter gradients, the scores and the visibility into fused [16,P] (column blocks 3,3,1,2,4 | 2 | 1) for
 * ONE all-reduce(sum), and SB200_UPDATE_DP_POST does the bookkeeping and Adam from the reduced buffer. */
enum { SB200_UPDATE_LOCAL = 0, SB200_UPDATE_DP_PRE = 1, SB200_UPDATE_DP_POST = 2 };
int sb200_surfel_update(int P, int mode, float* const* params, float* const* exp_avg, float* const* exp_avg_sq,
                        const float* d_means3D, const float* d_colors, const float* d_opacity, const float* d_scales,
                        const float* d_rotations, const float* cur_scores, const int* radii,
                        const uint8_t* stable_mask, float* local_scores, float* global_scores, float* kf_max_scores,
                        int64_t* kf_id, int64_t kf, float* fused, const float* lr, float b1, float b2, float eps,
                        void* stream);

/* sb200_backward_update: sb200_backward and sb200_surfel_update as ONE call for the fused mapping step - the backward
 * blend, then a single per-surfel pass that leaves a surfel the view did not see (radii <= 0) untouched and takes a
 * visible one from its accumulation record through the backward preprocess (cuda_rasterizer/backward.cu:849-1045), the
 * activations' backward, add_records, weighting_grad and the masked Adam step without the dense [P,*] gradient arrays
 * in between (72 B written and ~60 B read back per surfel by the two-call form).
 *   means3D / scales / rotations: the rasterizer inputs of this view (activated); params: the five RAW parameter arrays
 *   (params[0] is means3D); the rest as sb200_surfel_update.
 *   mode 0: single-GPU update (SB200_UPDATE_LOCAL semantics).
 *   mode 1: multi-GPU stage - does what sb200_dp_stage does (stage rows of visible surfels, visibility bytes, signal to
 *           the peers) on the exchange handle `dp` and updates nothing; exp_avg .. lr may be NULL.
 *   mode | SB200_BWD_RECORDS_ZERO: grad_ws is known to be all zero on entry, skip its memset.  The call always hands
 *           grad_ws back all zero (the per-surfel pass zeroes the records it consumed; only visible surfels ever
 *           receive any), so a caller that uses nothing else on that buffer sets the bit on every call but the first.
 *   R < 0 (test hook): skip the blend, grad_ws already holds the view's accumulation records (left untouched). */
#define SB200_BWD_RECORDS_ZERO 4
int sb200_backward_update(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                          const float* means3D, const float* scales, const float* rotations,
                          const float* viewmatrix, const float* projmatrix, const int* radii,
                          const void* geom_ws, const uint32_t* point_list, const void* image_ws,
                          const float* out_color, const float* out_others, const float* dL_dout_color,
                          const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws, int mode,
                          float* const* params, float* const* exp_avg, float* const* exp_avg_sq,
                          const uint8_t* stable_mask, float* local_scores, float* global_scores,
                          float* kf_max_scores, int64_t* kf_id, int64_t kf, const float* lr, float b1,
                          float b2, float eps, void* dp, void* stream);

/* ---- multi-GPU exchange of the mapping step over NVLink peer memory ----------------------
 * One optimizer step takes ONE view per rank (window views partitioned over the GPUs of a box, surfels replicated).
 * The calls below replace "sb200_surfel_update(DP_PRE) -> NCCL all-reduce of [16,P] -> sb200_surfel_update(DP_POST)" by
 * three kernels that read their peers' buffers directly and hand over to each other through flags in those buffers
 * (system-scope release / acquire; no barrier kernel, no collective call).  Ownership is cyclic over 256-surfel
 * chunks: surfel i is owned by rank (i / 256) % n_ranks.
 *   sb200_dp_create : rank's handle.  peer_bases: HOST array of n_ranks (<= 16) DEVICE pointers, entry q = rank q's
 *                     exchange buffer of sb200_dp_buffer_bytes(capacity, n_ranks) bytes, zero-filled, mapped into this
 *                     process (e.g. one torch symmetric-memory allocation per rank); every rank calls the three step
 *                     functions in lockstep, once per optimizer step, each on its own stream.
 *   sb200_dp_stage  : backward of the activations; for a visible surfel (radii > 0) writes its 13 raw-parameter
 *                     gradients and two scores to the rank's stage row (16 floats) and its visibility byte; the last
 *                     CTA signals the peers.  (sb200_backward_update with mode 1 does the same straight from the
 *                     accumulation records.)
 *   sb200_dp_reduce : waits for every peer's stage; for the owned surfels loads the stage rows of every rank that saw
 *                     them, applies weighting_grad per view (scripts/gaussian/gaussian_utils.py:56-62) and folds the
 *                     views in rank order: gradient sum, score-0 sum, score-1 max, score-0 max -> the owner's result row
 *                     (16 floats); the rank of the first arg-max of score 0 rides in the touched byte; signals.
 *   sb200_dp_post   : waits for every peer's reduce; every rank, every surfel some view saw: loads the owner's result
 *                     row, applies add_records (scripts/gaussian/gaussian_model.py:163-168) for the folded views (keyframe
 *                     id of the arg-max view = kf_ids[rank], a HOST array with one id per rank) and the
 *                     masked Adam step (cuda_rasterizer/adam.cu:9-38) to the local replica.  All ranks read the same
 *                     bytes: the replicas stay bit-identical, and equal a single process taking the views in rank order.
 *   sb200_dp_error  : non-zero if a wait ever gave up (a peer did not signal within 20 s); synchronises the stream.
 * params / exp_avg / exp_avg_sq / lr: as sb200_surfel_update. */
size_t sb200_dp_buffer_bytes(int64_t capacity, int n_ranks);
void* sb200_dp_create(int rank, int n_ranks, int64_t capacity, const void* const* peer_bases);
void sb200_dp_destroy(void* handle);
int sb200_dp_slice(void* handle, int P);
int sb200_dp_error(void* handle, void* stream);
/* diagnostics: nanoseconds the first CTA of this rank's reduce ([0]) and post ([1]) kernels spent waiting for the
 * slowest peer since the last call (i.e. rank skew, not work); resets the counters; synchronises the stream. */
int sb200_dp_wait_ns(void* handle, uint64_t out[2], void* stream);
int sb200_dp_stage(void* handle, int P, const float* d_means3D, const float* d_colors, const float* d_opacity,
                   const float* d_scales, const float* d_rotations, const float* cur_scores, const int* radii,
                   const float* opacity_raw, const float* scaling_raw, const float* rotation_raw, void* stream);
int sb200_dp_reduce(void* handle, int P, const float* global_scores, void* stream);
int sb200_dp_post(void* handle, int P, const int64_t* kf_ids, float* const* params, float* const* exp_avg, float* const* exp_avg_sq,
                  const uint8_t* stable_mask, float* local_scores, float* global_scores, float* kf_max_scores,
                  int64_t* kf_id, const float* lr, float b1, float b2, float eps, void* stream);

/* ---- surfel add / prune churn --------------------------------------------
 * sb200_knn_mean_dist2: out[i] = mean of the squared distances from points[i] to its 3 nearest OTHER points, i.e.
 * (KDTree(points).query(points, k=4)[0][:, 1:] ** 2).mean(1) of the reference's distCUDA2
 * (scripts/gaussian/gaussian_utils.py:12-16), computed on the device without the host round trip.  points [N,3].
 * Fewer than 4 points give +inf, like the KD-tree's missing neighbours. */
int sb200_knn_mean_dist2(int N, const float* points, float* out, void* stream);

/* sb200_prune_rows: stable compaction of up to SB200_PRUNE_MAX_ARRAYS per-surfel arrays by one mask, i.e.
 * x = x[~delete_mask] for the five parameters, their ten Adam moments and the score bookkeeping at once
 * (prune_tensors_from_optimizer + update_records("prune"), scripts/gaussian/gaussian_base.py:96-112).
 *   delete_mask: bool [P], 1 = drop the row;  src / dst / row_bytes: HOST arrays of n_arrays device pointers / row
 *   sizes in bytes; dst[a] must not alias src[a] (capacity buffers are used in pairs);
 *   scratch: sb200_prune_scratch_bytes(P) bytes;  *kept_host receives the surviving row count (the only host
 *   synchronisation; the reference synchronises once per indexed tensor). */
#define SB200_PRUNE_MAX_ARRAYS 32
size_t sb200_prune_scratch_bytes(int64_t P);
int sb200_prune_rows(int P, const uint8_t* delete_mask, int n_arrays, const void* const* src, void* const* dst,
                     const int* row_bytes, void* scratch, int* kept_host, void* stream);

/* sb200_prune_rows_static: sb200_prune_rows for callers that cannot allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same compaction, same kernels, same dst contents; the differences:
 *   - no scratch argument: the block counts live in a file-scope `static __device__` array of the library
 *     (SB200_PRUNE_STATIC_MAX_ROWS / 1024 + 2 words = 64 KiB + 8 B, held by every GPU that loads the library until
 *     exit).  P > SB200_PRUNE_STATIC_MAX_ROWS is rejected with return 1 and a message in sb200_last_error();
 *   - kept_dev is a DEVICE pointer to one int32 that receives the surviving row count in stream order; there is
 *     no host read-back and no synchronisation (P == 0 stores 0);
 *   - the call makes no cudaMalloc / cudaFree / allocating library call: three launches and one 4-byte
 *     device-to-device copy on `stream` (the first call on each device also resolves the array's address with
 *     cudaGetSymbolAddress); it is not bracketed by the optional stage profiling, whose events are allocations.
 * Every call on a device shares the one array, which keeps the previous call's contents; nothing in it has to start
 * at zero (each call rewrites every word it reads before reading it).  Calls that overlap on the same GPU (two
 * streams, two threads) RACE on it: serialise them, or use sb200_prune_rows with one scratch buffer per stream. */
#define SB200_PRUNE_STATIC_MAX_ROWS (16 * 1024 * 1024)
int sb200_prune_rows_static(int P, const uint8_t* delete_mask, int n_arrays, const void* const* src, void* const* dst,
                            const int* row_bytes, int* kept_dev, void* stream);

/* ---- precomputed splat transforms (the rasterizer's `cov3D_precomp` = transMat_precomp input; unused by VINGS-Mono) ----
 * transMat_precomp [P,9]: rows Tu, Tv, Tw of every surfel's splat-to-pixel transform, used instead of scales +
 * rotations (cuda_rasterizer/forward.cu:204-219: the normal becomes (0,0,1)).  Stage 2 is sb200_forward_finish.
 * The backward (backward.cu:875-883,950-964) then produces dL_dtransMat [P,9] (required), dL_dcolors, dL_dopacity,
 * scores and the dL_dmeans2D heuristic; dL_dmeans3D / dL_dscales / dL_drotations are written as zero rows. */
int sb200_forward_begin_precomp(int P, int W, int H, const float* means3D, const float* transMat_precomp,
                                const float* opacities, const float* colors, const float* viewmatrix,
                                const float* projmatrix, void* geom_ws, int* radii, int64_t* num_rendered_host,
                                void* stream);
int sb200_backward_precomp(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                           const float* transMat_precomp, const float* viewmatrix, const float* projmatrix,
                           const int* radii, const void* geom_ws, const uint32_t* point_list, const void* image_ws,
                           const float* out_color, const float* out_others, const float* dL_dout_color,
                           const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws, float* dL_dmeans3D,
                           float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dscales,
                           float* dL_drotations, float* dL_dtransMat, float* scores, void* stream);

/* ---- spherical-harmonics colours (the rasterizer's `shs` / `sh_degree` inputs; unused by VINGS-Mono) ----------
 * The reference evaluates SH inside its preprocess kernels (computeColorFromSH, cuda_rasterizer/forward.cu:28-79,
 * backward.cu:20-139); here it is a stage in front of / behind the rasterizer:
 *   sb200_sh_forward : colors[P,3] = max(0, sum_k basis_k(dir) * shs[:,k,:] + 0.5), dir = normalize(means3D - campos),
 *                      degree 0..3, shs [P, n_coeffs, 3] with n_coeffs >= (degree+1)^2; clamped[P,3] (u8) = was < 0.
 *                      Pass `colors` to sb200_forward_begin as the colours.
 *   sb200_sh_backward: from sb200_backward's dL_dcolors: dL_dsh [P, n_coeffs, 3] (fully written; zero rows for
 *                      radii <= 0, zero above the active degree) and dL_dmeans3D += the view-direction term. */
int sb200_sh_forward(int P, int degree, int n_coeffs, const float* means3D, const float* campos /* [3] */,
                     const float* shs, float* colors, uint8_t* clamped, void* stream);
int sb200_sh_backward(int P, int degree, int n_coeffs, const float* means3D, const float* campos, const float* shs,
                      const uint8_t* clamped, const int* radii, const float* dL_dcolors, float* dL_dmeans3D,
                      float* dL_dsh, void* stream);

/* ---- building blocks exposed for parity tests ----------------------------------------------- */
/* Stable LSD radix sort of n (u32 key, u32 value) pairs on key bits [0,end_bit), the sort both binning sorts
 * run on.  scratch = sb200_binning_bytes(n) bytes whose tile ids A / surfel ids A arrays (see layout) hold the
 * input keys / values and are clobbered.  vals_out (caller buffer of n) receives the values in sorted order.
 * mode bit 0: 2048-pair tiles instead of 4096.  mode bit 1: the surfel-sort form - values are the indices 0..n-1,
 * keys equal to 0xFFFFFFFF are dropped, *n_valid_dev (device u32) receives the number of pairs kept. */
int sb200_sort_pairs(void* binning_ws, int64_t n, int end_bit, int mode, uint32_t* vals_out, uint32_t* n_valid_dev,
                     void* stream);

/* Test hook for reference quirk Q17: the reference's forward stores a per-tile `max_contrib`
 * (cuda_rasterizer/forward.cu:496-502) that its backward uses to drop whole 32-surfel buckets (backward.cu:503-505),
 * and that value is not the tile maximum (cub::BlockReduce<.,256> on a 16x16 block).  With per_tile_cap (device
 * u32[tiles], must stay alive) set, the process's following sb200_backward calls stop every tile's walk at
 * ceil(cap/32) buckets exactly like the reference does, which lets a test show that this quirk explains ALL of the
 * gradient difference on the surfels it excludes.  NULL switches it off (the default). */
int sb200_debug_walk_cap(const uint32_t* per_tile_cap);

#ifdef __cplusplus
}
#endif
#endif
