// C ABI of the surfel rasterizer (declared in include/surfel_b200.h): stage sequencing, workspace
// carving and error reporting.  Replaces the host drivers CudaRasterizer::Rasterizer::forward/backward
// (DSR/cuda_rasterizer/rasterizer_impl.cu:305-478, 482-599).
#include "common.cuh"
#include "../../include/surfel_b200.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <vector>

static thread_local char g_err[512] = "";
#ifndef SB_PAIR_SMALL
#define SB_PAIR_SMALL 0  // pair sort on 4096-pair tiles (1 = 2048)
#endif

// ---- optional per-stage device timing (CUDA events on the launching stream) and launch counting ----
static std::atomic<long long> g_launches{0};
void sb_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

struct StageRec { int stage; cudaEvent_t e0, e1; };
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static std::vector<StageRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;

static cudaEvent_t prof_event() {
    cudaEvent_t e;
    if (!g_prof_pool.empty()) { e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEventCreate(&e);
    return e;
}

SbStageScope::SbStageScope(int stage_, cudaStream_t s_) : on(false), stage(stage_), s(s_), e0(nullptr) {
    std::lock_guard<std::mutex> g(g_prof_mu);
    if (!g_prof_on) return;
    on = true;
    e0 = prof_event();
    cudaEventRecord(e0, s);
}
SbStageScope::~SbStageScope() {
    if (!on) return;
    std::lock_guard<std::mutex> g(g_prof_mu);
    cudaEvent_t e1 = prof_event();
    cudaEventRecord(e1, s);
    g_prof_recs.push_back({stage, e0, e1});
}
typedef SbStageScope StageScope;

void sb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" {

int sb200_abi_version(void) { return 7; }

long long sb200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int sb200_profile_enable(int on) {
    std::lock_guard<std::mutex> g(g_prof_mu);
    g_prof_on = on != 0;
    return 0;
}

int sb200_profile_read(double* ms, long long* count) {
    std::lock_guard<std::mutex> g(g_prof_mu);
    for (int i = 0; i < SB200_NUM_STAGES; i++) { ms[i] = 0.0; count[i] = 0; }
    for (StageRec& r : g_prof_recs) {
        float t = 0.f;
        cudaError_t e = cudaEventSynchronize(r.e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, r.e0, r.e1);
        if (e != cudaSuccess) { sb_set_error("profile_read: %s", cudaGetErrorString(e)); return 2; }
        ms[r.stage] += (double)t;
        count[r.stage] += 1;
        g_prof_pool.push_back(r.e0);
        g_prof_pool.push_back(r.e1);
    }
    g_prof_recs.clear();
    return 0;
}
const char* sb200_last_error(void) { return g_err; }

size_t sb200_geom_bytes(int64_t P) { return sb_geom_layout(P).total; }
size_t sb200_image_bytes(int W, int H) { return sb_image_layout(W, H).total; }
size_t sb200_binning_bytes(int64_t R) { return sb_bin_layout(R).total; }
size_t sb200_pair_list_bytes(int64_t R, int W, int H) { return sb_pair_list_layout(R, W, H).total; }
size_t sb200_grad_bytes(int64_t P) { return sb_align((size_t)(P > 0 ? P : 0) * SB_GRAD_F * 4); }

void sb200_geom_layout(int64_t P, size_t out[5]) {
    SbGeomLayout L = sb_geom_layout(P);
    out[0] = L.rec; out[1] = L.depth; out[2] = L.rect; out[3] = L.tiles; out[4] = L.sort_v[1];
}
void sb200_image_layout(int W, int H, size_t out[3]) {
    SbImageLayout L = sb_image_layout(W, H);
    out[0] = L.ranges; out[1] = L.final_T; out[2] = L.n_contrib;
}
void sb200_binning_layout(int64_t R, size_t out[4]) {
    SbBinLayout L = sb_bin_layout(R);
    out[0] = L.keys_a; out[1] = L.keys_b; out[2] = L.vals_a; out[3] = L.vals_b;
}

// One event per (thread, device): lets forward_begin wait for the read-back of R alone while the surfel sort,
// queued behind it on the same stream, keeps the GPU busy.
static cudaEvent_t sync_event() {
    thread_local cudaEvent_t ev[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (!ev[dev] && cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming) != cudaSuccess) ev[dev] = nullptr;
    return ev[dev];
}

static int check_dims(int P, int W, int H) {
    if (P < 0 || W <= 0 || H <= 0) { sb_set_error("bad dimensions P=%d W=%d H=%d", P, W, H); return 1; }
    if ((W + SB_TILE_X - 1) / SB_TILE_X > 65535 || (H + SB_TILE_Y - 1) / SB_TILE_Y > 65535) {
        sb_set_error("image %dx%d exceeds the 65535-tile rect encoding", W, H);
        return 1;
    }
    return 0;
}

static int forward_begin_impl(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                              const float* transmat, const float* opacities, const float* colors,
                              const float* viewmatrix, const float* projmatrix, void* geom_ws, int* radii,
                              int64_t* num_rendered_host, void* stream) {
    g_err[0] = 0;
    if (check_dims(P, W, H)) return 1;
    if (!num_rendered_host) { sb_set_error("num_rendered_host is NULL"); return 1; }
    *num_rendered_host = 0;
    if (P == 0) return 0;  // reference: rasterize_points.cu:106
    if (!means3D || (!transmat && (!scales || !rotations)) || !opacities || !colors || !viewmatrix || !projmatrix ||
        !geom_ws || !radii) {
        sb_set_error("forward_begin: NULL input");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    {
        StageScope sc(SB200_STAGE_PREPROCESS, s);
        rc = sb_launch_preprocess(P, W, H, means3D, scales, rotations, transmat, opacities, colors, viewmatrix,
                                  projmatrix, (char*)geom_ws, radii, s);
    }
    if (rc) return rc;
    SbGeomLayout L = sb_geom_layout(P);
    cudaEvent_t ev = sync_event();
    SB_CUDA_CHECK(cudaMemcpyAsync(num_rendered_host, (char*)geom_ws + L.counts, 8, cudaMemcpyDeviceToHost, s));
    if (ev) SB_CUDA_CHECK(cudaEventRecord(ev, s));
    {
        StageScope sc(SB200_STAGE_SURFEL_SORT, s);
        rc = sb_launch_surfel_sort(P, (char*)geom_ws, s);
    }
    if (rc) return rc;
    if (ev) SB_CUDA_CHECK(cudaEventSynchronize(ev));
    else SB_CUDA_CHECK(cudaStreamSynchronize(s));
    return 0;
}

int sb200_forward_begin(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                        const float* opacities, const float* colors, const float* viewmatrix, const float* projmatrix,
                        void* geom_ws, int* radii, int64_t* num_rendered_host, void* stream) {
    return forward_begin_impl(P, W, H, means3D, scales, rotations, nullptr, opacities, colors, viewmatrix, projmatrix,
                              geom_ws, radii, num_rendered_host, stream);
}

int sb200_forward_begin_precomp(int P, int W, int H, const float* means3D, const float* transMat_precomp,
                                const float* opacities, const float* colors, const float* viewmatrix,
                                const float* projmatrix, void* geom_ws, int* radii, int64_t* num_rendered_host,
                                void* stream) {
    // (a NULL transMat_precomp fails forward_begin_impl's input check: neither a transform nor scales + rotations)
    return forward_begin_impl(P, W, H, means3D, nullptr, nullptr, transMat_precomp, opacities, colors, viewmatrix,
                              projmatrix, geom_ws, radii, num_rendered_host, stream);
}

static int forward_finish_impl(int P, int W, int H, int64_t R, int R_is_capacity, const float* bg, void* geom_ws,
                               void* binning_ws, uint32_t* point_list, void* image_ws, float* out_color,
                               float* out_others, cudaStream_t s) {
    if (check_dims(P, W, H)) return 1;
    if (!bg || !image_ws || !out_color || !out_others) { sb_set_error("forward_finish: NULL input"); return 1; }
    if (R > 0 && (!geom_ws || !binning_ws || !point_list)) { sb_set_error("forward_finish: NULL workspace"); return 1; }
    const int tiles = ((W + SB_TILE_X - 1) / SB_TILE_X) * ((H + SB_TILE_Y - 1) / SB_TILE_Y);
    SbImageLayout IL = sb_image_layout(W, H);
    uint32_t* tile_counts = (uint32_t*)((char*)image_ws + IL.counts);
    uint2* ranges = (uint2*)((char*)image_ws + IL.ranges);
    uint32_t* order = (uint32_t*)((char*)image_ws + IL.order);
    const SbSortPlan pl = sb_sort_plan((int)sb_higher_msb((uint32_t)tiles));  // the reference's tile bit count (:410)
    int rc;
    SB_CUDA_CHECK(cudaMemsetAsync(tile_counts, 0, (size_t)tiles * 4, s));
    if (R > 0 && P > 0) {
        SbBinLayout BL = sb_bin_layout(R);
        SbGeomLayout GL = sb_geom_layout(P);
        char* bw = (char*)binning_ws;
        uint32_t* kbuf[2] = {(uint32_t*)(bw + BL.keys_b), (uint32_t*)(bw + BL.keys_a)};
        uint32_t* vbuf[2] = {(uint32_t*)(bw + BL.vals_b), (uint32_t*)(bw + BL.vals_a)};
        SB_CUDA_CHECK(cudaMemsetAsync(bw + BL.zero_begin, 0,
                                      (BL.status - BL.zero_begin) + sb_sort_status_bytes(R, pl.passes, SB_PAIR_SMALL), s));
        {
            StageScope sc(SB200_STAGE_DUPLICATE, s);
            rc = sb_launch_duplicate(P, W, H, R, R_is_capacity, (const char*)geom_ws, (uint32_t*)(bw + BL.chunk_first),
                                     tile_counts, kbuf[1], vbuf[1], s);
        }
        if (rc) return rc;
        {
            StageScope sc(SB200_STAGE_RANGES, s);
            rc = sb_launch_tile_scan(tiles, pl, tile_counts, ranges, order, (uint32_t*)(bw + BL.hist), s);
        }
        if (rc) return rc;
        {
            StageScope sc(SB200_STAGE_SORT, s);
            // little-endian: the low word of the 64-bit pair count (R < 2^30 is checked on the capacity)
            const uint32_t* n_dev = R_is_capacity ? (const uint32_t*)((const char*)geom_ws + GL.counts) : nullptr;
            rc = sb_launch_sort_passes(pl, SB_PAIR_SMALL, kbuf[1], vbuf[1], kbuf, vbuf, point_list, R, n_dev,
                                       (const uint32_t*)(bw + BL.hist), (uint32_t*)(bw + BL.tickets),
                                       (uint32_t*)(bw + BL.status), s);
        }
        if (rc) return rc;
    } else {
        StageScope sc(SB200_STAGE_RANGES, s);
        rc = sb_launch_tile_scan(tiles, pl, tile_counts, ranges, order, nullptr, s);
        if (rc) return rc;
    }
    // with R == 0 every range is (0,0): the blend writes background / zeros, as the reference does
    StageScope sc(SB200_STAGE_BLEND_FWD, s);
    uint32_t* hit_masks = (uint32_t*)((char*)point_list + sb_pair_list_layout(R, W, H).masks);
    return sb_launch_blend_forward(W, H, bg, (const char*)geom_ws, point_list, (char*)image_ws, out_color, out_others,
                                   hit_masks, s);
}

int sb200_forward_finish(int P, int W, int H, int64_t R, const float* bg, void* geom_ws, void* binning_ws,
                         uint32_t* point_list, void* image_ws, float* out_color, float* out_others, void* stream) {
    g_err[0] = 0;
    return forward_finish_impl(P, W, H, R, 0, bg, geom_ws, binning_ws, point_list, image_ws, out_color, out_others,
                               (cudaStream_t)stream);
}

int sb200_forward_async(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                        const float* opacities, const float* colors, const float* viewmatrix, const float* projmatrix,
                        const float* bg, void* geom_ws, int* radii, void* binning_ws, int64_t pair_capacity,
                        uint32_t* point_list, void* image_ws, float* out_color, float* out_others,
                        int64_t* num_rendered_host, void* sync, void* stream) {
    g_err[0] = 0;
    if (check_dims(P, W, H)) return 1;
    if (!num_rendered_host) { sb_set_error("num_rendered_host is NULL"); return 1; }
    if (pair_capacity < 1) { sb_set_error("forward_async: pair_capacity must be positive"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    cudaEvent_t ev = (cudaEvent_t)sync;
    if (!ev) { sb_set_error("forward_async: sync handle is NULL (sb200_sync_create)"); return 1; }
    *num_rendered_host = 0;
    int rc;
    if (P > 0) {
        if (!means3D || !scales || !rotations || !opacities || !colors || !viewmatrix || !projmatrix || !geom_ws || !radii) {
            sb_set_error("forward_async: NULL input");
            return 1;
        }
        {
            StageScope sc(SB200_STAGE_PREPROCESS, s);
            rc = sb_launch_preprocess(P, W, H, means3D, scales, rotations, nullptr, opacities, colors, viewmatrix,
                                      projmatrix, (char*)geom_ws, radii, s);
        }
        if (rc) return rc;
        SbGeomLayout L = sb_geom_layout(P);
        SB_CUDA_CHECK(cudaMemcpyAsync(num_rendered_host, (char*)geom_ws + L.counts, 8, cudaMemcpyDeviceToHost, s));
        SB_CUDA_CHECK(cudaEventRecord(ev, s));
        StageScope sc(SB200_STAGE_SURFEL_SORT, s);
        rc = sb_launch_surfel_sort(P, (char*)geom_ws, s);
        if (rc) return rc;
    } else {
        SB_CUDA_CHECK(cudaEventRecord(ev, s));
    }
    return forward_finish_impl(P, W, H, P > 0 ? pair_capacity : 0, 1, bg, geom_ws, binning_ws, point_list, image_ws,
                               out_color, out_others, s);
}

int sb200_forward_wait(void* sync) {
    g_err[0] = 0;
    if (!sync) { sb_set_error("forward_wait: sync handle is NULL"); return 1; }
    SB_CUDA_CHECK(cudaEventSynchronize((cudaEvent_t)sync));
    return 0;
}

void* sb200_sync_create(void) {
    g_err[0] = 0;
    cudaEvent_t ev = nullptr;
    cudaError_t e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    if (e != cudaSuccess) { sb_set_error("sync_create: %s", cudaGetErrorString(e)); return nullptr; }
    return (void*)ev;
}

void sb200_sync_destroy(void* sync) {
    if (sync) cudaEventDestroy((cudaEvent_t)sync);
}

// test hook: per-tile cap of the backward walk for the process's next backward calls (NULL = off).  Process-wide, not
// thread-local: torch runs an autograd backward on its engine's worker thread, not on the thread that set the hook.
static std::atomic<const uint32_t*> g_walk_cap{nullptr};

static int backward_impl(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                   const float* means3D, const float* scales, const float* rotations, const float* transmat,
                   const float* viewmatrix,
                   const float* projmatrix, const int* radii, const void* geom_ws, const uint32_t* point_list,
                   const void* image_ws, const float* out_color, const float* out_others,
                   const float* dL_dout_color, const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws,
                   float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dscales,
                   float* dL_drotations, float* dL_dtransMat, float* scores, void* stream) {
    g_err[0] = 0;
    if (check_dims(P, W, H)) return 1;
    if (P == 0) return 0;  // reference: rasterize_points.cu:217
    if ((transmat ? !dL_dtransMat : (!means3D || !scales || !rotations)) ||
        !geom_ws || !image_ws || !grad_ws || !radii || !dL_dout_color || !dL_dout_others || !pixel_mask ||
        !dL_dmeans3D || !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dscales || !dL_drotations || !scores) {
        sb_set_error("backward: NULL input");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (R > 0) {
        StageScope sc(SB200_STAGE_BLEND_BWD, s);
        SB_CUDA_CHECK(cudaMemsetAsync(grad_ws, 0, (size_t)P * SB_GRAD_F * 4, s));
        rc = sb_launch_blend_backward(W, H, bg, (const char*)geom_ws, point_list, (const char*)image_ws, out_color,
                                      out_others, dL_dout_color, dL_dout_others, (const bool*)pixel_mask,
                                      (float*)grad_ws,
                                      (const uint32_t*)((const char*)point_list + sb_pair_list_layout(R, W, H).masks),
                                      g_walk_cap.load(std::memory_order_acquire), s);
        if (rc) return rc;
    } else {
        SB_CUDA_CHECK(cudaMemsetAsync(grad_ws, 0, (size_t)P * SB_GRAD_F * 4, s));
    }
    StageScope sc(SB200_STAGE_PREPROCESS_BWD, s);
    return sb_launch_preprocess_backward(P, W, H, tanfovx, tanfovy, means3D, scales, rotations, transmat, viewmatrix, projmatrix,
                                         radii, (const char*)geom_ws, (const float*)grad_ws, dL_dmeans3D, dL_dmeans2D,
                                         dL_dcolors, dL_dopacity, dL_dscales, dL_drotations, dL_dtransMat, scores, s);
}

int sb200_backward(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                   const float* means3D, const float* scales, const float* rotations, const float* viewmatrix,
                   const float* projmatrix, const int* radii, const void* geom_ws, const uint32_t* point_list,
                   const void* image_ws, const float* out_color, const float* out_others,
                   const float* dL_dout_color, const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws,
                   float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dscales,
                   float* dL_drotations, float* dL_dtransMat, float* scores, void* stream) {
    return backward_impl(P, W, H, R, bg, tanfovx, tanfovy, means3D, scales, rotations, nullptr, viewmatrix, projmatrix,
                         radii, geom_ws, point_list, image_ws, out_color, out_others, dL_dout_color, dL_dout_others,
                         pixel_mask, grad_ws, dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dscales,
                         dL_drotations, dL_dtransMat, scores, stream);
}

int sb200_backward_precomp(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                           const float* transMat_precomp, const float* viewmatrix, const float* projmatrix,
                           const int* radii, const void* geom_ws, const uint32_t* point_list, const void* image_ws,
                           const float* out_color, const float* out_others, const float* dL_dout_color,
                           const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws, float* dL_dmeans3D,
                           float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dscales,
                           float* dL_drotations, float* dL_dtransMat, float* scores, void* stream) {
    return backward_impl(P, W, H, R, bg, tanfovx, tanfovy, nullptr, nullptr, nullptr, transMat_precomp, viewmatrix,
                         projmatrix, radii, geom_ws, point_list, image_ws, out_color, out_others, dL_dout_color,
                         dL_dout_others, pixel_mask, grad_ws, dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dopacity,
                         dL_dscales, dL_drotations, dL_dtransMat, scores, stream);
}

int sb200_debug_walk_cap(const uint32_t* per_tile_cap) {
    g_walk_cap.store(per_tile_cap, std::memory_order_release);
    return 0;
}

int sb200_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, void* stream) {
    g_err[0] = 0;
    (void)projmatrix;  // the reference computes p_hom from it and never uses the result (auxiliary.h:207-212)
    if (P < 0) { sb_set_error("bad P"); return 1; }
    return sb_launch_mark_visible(P, means3D, viewmatrix, (bool*)present, (cudaStream_t)stream);
}

int sb200_adam_update(float* param, const float* param_grad, float* exp_avg, float* exp_avg_sq,
                      const uint8_t* visible, float lr, float b1, float b2, float eps, uint32_t N, uint32_t M,
                      void* stream) {
    g_err[0] = 0;
    StageScope sc(SB200_STAGE_ADAM, (cudaStream_t)stream);
    return sb_launch_adam(param, param_grad, exp_avg, exp_avg_sq, (const bool*)visible, lr, b1, b2, eps, N, M,
                          (cudaStream_t)stream);
}

size_t sb200_map_loss_bytes(int W, int H) { return sb_map_loss_bytes(W, H); }

int sb200_map_loss(int W, int H, const float* color, const float* allmap, const float* gt_rgb, const float* gt_depth,
                   const float* gt_depth_cov, const float* w2c, float cx, float cy, float fx, float fy, float w_rgb,
                   float w_normal, float w_alpha, float w_depth, float w_dist, void* ws, float* loss,
                   float* dL_dcolor, float* dL_dallmap, void* stream) {
    g_err[0] = 0;
    if (W <= 0 || H <= 0) { sb_set_error("map_loss: bad dimensions %dx%d", W, H); return 1; }
    if (!color || !allmap || !gt_rgb || !gt_depth || !gt_depth_cov || !w2c || !ws || !loss || !dL_dcolor || !dL_dallmap) {
        sb_set_error("map_loss: NULL input");
        return 1;
    }
    StageScope sc(SB200_STAGE_MAP_LOSS, (cudaStream_t)stream);
    return sb_launch_map_loss(W, H, color, allmap, gt_rgb, gt_depth, gt_depth_cov, w2c, cx, cy, fx, fy, w_rgb,
                              w_normal, w_alpha, w_depth, w_dist, (char*)ws, loss, dL_dcolor, dL_dallmap,
                              (cudaStream_t)stream);
}

int sb200_activate(int P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                   float* opacities, float* scales, float* rotations, void* stream) {
    g_err[0] = 0;
    if (P < 0) { sb_set_error("activate: bad P"); return 1; }
    if (P > 0 && (!opacity_raw || !scaling_raw || !rotation_raw || !opacities || !scales || !rotations)) {
        sb_set_error("activate: NULL input");
        return 1;
    }
    StageScope sc(SB200_STAGE_ACTIVATE, (cudaStream_t)stream);
    return sb_launch_activate(P, opacity_raw, scaling_raw, rotation_raw, opacities, scales, rotations,
                              (cudaStream_t)stream);
}

int sb200_sort_pairs(void* binning_ws, int64_t n, int end_bit, int mode, uint32_t* vals_out, uint32_t* n_valid_dev,
                     void* stream) {
    g_err[0] = 0;
    if (n <= 0) return 0;
    if (end_bit < 1 || end_bit > 32) { sb_set_error("sort_pairs: end_bit must be in [1,32]"); return 1; }
    const int small = mode & 1, drop = (mode >> 1) & 1;
    if (drop && !n_valid_dev) { sb_set_error("sort_pairs: drop mode needs n_valid_dev"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    SbBinLayout BL = sb_bin_layout(n);
    char* bw = (char*)binning_ws;
    const SbSortPlan pl = sb_sort_plan(end_bit);
    uint32_t* kbuf[2] = {(uint32_t*)(bw + BL.keys_b), (uint32_t*)(bw + BL.keys_a)};
    uint32_t* vbuf[2] = {(uint32_t*)(bw + BL.vals_b), (uint32_t*)(bw + BL.vals_a)};
    SB_CUDA_CHECK(cudaMemsetAsync(bw + BL.hist, 0, (BL.status - BL.hist) + sb_sort_status_bytes(n, pl.passes, small), s));
    int rc = sb_launch_sort_hist(kbuf[1], n, pl, drop, (uint32_t*)(bw + BL.hist), n_valid_dev, s);
    if (rc) return rc;
    return sb_launch_sort_passes(pl, small, kbuf[1], drop ? nullptr : vbuf[1], kbuf, vbuf, vals_out, n,
                                 drop ? n_valid_dev : nullptr, (const uint32_t*)(bw + BL.hist),
                                 (uint32_t*)(bw + BL.tickets), (uint32_t*)(bw + BL.status), s);
}

}  // extern "C"
