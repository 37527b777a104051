// Shared definitions for the B200 surfel rasterizer kernels (sm_100a).
//
// Domain vocabulary follows the reference (diff-surfel-rasterization, "DSR"):
//   surfel  = one 2D Gaussian, P of them
//   tile    = 16x16 pixel block (DSR/cuda_rasterizer/config.h:17-18) -- part of the key contract
//   pair    = one (tile, surfel) overlap, R = num_rendered of them
//   record  = the packed per-surfel render record this implementation blends from
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define SB_TILE_X 16
#define SB_TILE_Y 16
#define SB_TILE_PIX 256
#define SB_NEAR 0.2f    // DSR/cuda_rasterizer/auxiliary.h:37
#define SB_FAR 100.0f   // DSR/cuda_rasterizer/auxiliary.h:38
#define SB_FILTER_INV_SQUARE 2.0f  // auxiliary.h:39

// Per-surfel render record: 8 x float4 = 128 B = one cache line, fetched as one unit by the blend
// kernels.  Quads 0-2 feed the division-free cull test, 3-6 the exact reference arithmetic, 7 the
// warp-level bounding-box test.
//   q0 = (A.x,  A.y,  A.z,  xy.x)      p(pixel) = C + A*dx + B*dy with (dx,dy) = xy - pixel, where
//   q1 = (B.x,  B.y,  B.z,  xy.y)      p = (px*Tw - Tu) x (py*Tw - Tv) is affine in the pixel
//   q2 = (C.x,  C.y,  C.z,  tau)       tau = padded cutoff: alpha >= 1/255 needs min(rho3d,rho2d) <= tau
//   q3 = (Tu.x, Tu.y, Tu.z, opacity)
//   q4 = (Tv.x, Tv.y, Tv.z, r)
//   q5 = (Tw.x, Tw.y, Tw.z, g)
//   q6 = (n.x,  n.y,  n.z,  b)
//   q7 = (xmin, xmax, ymin, ymax)      pixel-space box outside of which the surfel cannot reach alpha >= 1/255
#define SB_REC_F4 8
#define SB_REC_CULL 0
#define SB_REC_TU 3
#define SB_REC_TV 4
#define SB_REC_TW 5
#define SB_REC_N 6
#define SB_REC_BBOX 7

// Per-surfel gradient accumulation record (floats), filled by the backward blend kernel with
// atomics and consumed by the backward preprocess kernel.
//   0..8  dL/dT (Tu, Tv, Tw)   9..10 dL/dmean2D (low-pass branch)   11..13 dL/dnormal
//   14    dL/dopacity          15..17 dL/dcolor                     18..19 scores
#define SB_GRAD_F 20

static inline __host__ __device__ size_t sb_align(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Radix sort geometry (onesweep).
#define SB_SORT_THREADS 256
#define SB_SORT_ITEMS 16        // pair sort: 4096-pair tiles
#define SB_SORT_ITEMS_SMALL 8   // surfel sort: 2048-surfel tiles (more CTAs for the shorter input)
#define SB_RADIX_BITS 8
#define SB_RADIX 256
#define SB_MAX_PASSES 4
#define SB_KEY_INVALID 0xffffffffu  // depth key of a surfel that touches no tile: dropped by the surfel sort
#define SB_DEPTH_BITS 32
#define SB_EMIT_CHUNK 1024  // pairs per work unit of the duplicate kernel

struct SbSortPlan { int end_bit, passes, width; };
SbSortPlan sb_sort_plan(int end_bit);

// Geometry workspace (caller-allocated, P-sized). Offsets in bytes.
struct SbGeomLayout {
    size_t rec;        // float4[8P]
    size_t depth;      // float[P]       view-space z of the visible surfels
    size_t rect;       // ushort4[P]     tile rect (min.x, min.y, max.x, max.y)
    size_t tiles;      // uint32[P]      tiles_touched
    size_t sort_k[2];  // uint32[P] x 2  depth keys: [1] is written by preprocess (SB_KEY_INVALID = not visible)
    size_t sort_v[2];  // uint32[P] x 2  surfel ids; after the 4 passes [1] holds the V visible ids in depth order
    size_t state;      // zeroed before every forward:
    size_t hist;       //   uint32[4*256]  digit bases of the surfel sort
    size_t tickets;    //   uint32[8]      [0..3] sort pass tickets, [4] duplicate ticket
    size_t counts;     //   uint64[2]      [0] R = sum of tiles_touched, [1] (low word) V = visible surfels
    size_t dup_state;  //   uint64[ceil(P/256)]  look-back words of the pair-offset scan
    size_t status;     //   uint32[4 * ceil(P/2048) * 256]  look-back words of the surfel sort
    size_t state_bytes;
    size_t total;
};

static inline __host__ SbGeomLayout sb_geom_layout(int64_t P) {
    SbGeomLayout L;
    size_t o = 0;
    size_t p = (size_t)(P > 0 ? P : 0);
    L.rec = o;        o = sb_align(o + p * SB_REC_F4 * 16);
    L.depth = o;      o = sb_align(o + p * 4);
    L.rect = o;       o = sb_align(o + p * 8);
    L.tiles = o;      o = sb_align(o + p * 4);
    for (int i = 0; i < 2; i++) { L.sort_k[i] = o; o = sb_align(o + p * 4); }
    for (int i = 0; i < 2; i++) { L.sort_v[i] = o; o = sb_align(o + p * 4); }
    L.state = o;
    L.hist = o;       o = sb_align(o + SB_MAX_PASSES * SB_RADIX * 4);
    L.tickets = o;    o = sb_align(o + 8 * 4);
    L.counts = o;     o = sb_align(o + 2 * 8);
    L.dup_state = o;  o = sb_align(o + ((p + 255) / 256 + 1) * 8);
    size_t nt = (p + SB_SORT_THREADS * SB_SORT_ITEMS_SMALL - 1) / (SB_SORT_THREADS * SB_SORT_ITEMS_SMALL);
    L.status = o;     o = sb_align(o + (size_t)SB_MAX_PASSES * nt * SB_RADIX * 4);
    L.state_bytes = o - L.state;
    L.total = o;
    return L;
}

// Image workspace (caller-allocated): tile ranges + the per-pixel state the backward needs.
struct SbImageLayout {
    size_t ranges;    // uint2[tiles]
    size_t final_T;   // float[3*HW]   T, M1, M2   (reference: ImageState::accum_alpha)
    size_t n_contrib; // uint32[2*HW]  last contributor, median contributor
    size_t order;     // uint32[tiles] launch order of the tiles, longest list first
    size_t counts;    // uint32[tiles] pairs per tile, accumulated by the duplicate kernel
    size_t total;
};

static inline __host__ SbImageLayout sb_image_layout(int W, int H) {
    SbImageLayout L;
    size_t tiles = (size_t)((W + SB_TILE_X - 1) / SB_TILE_X) * ((H + SB_TILE_Y - 1) / SB_TILE_Y);
    size_t hw = (size_t)W * H;
    size_t o = 0;
    L.ranges = o;    o = sb_align(o + tiles * 8);
    L.final_T = o;   o = sb_align(o + hw * 12);
    L.n_contrib = o; o = sb_align(o + hw * 8);
    L.order = o;     o = sb_align(o + tiles * 4);
    L.counts = o;    o = sb_align(o + tiles * 4);
    L.total = o;
    return L;
}

// Pair list (caller-allocated, persists forward -> backward): the sorted surfel ids, then the forward's hit masks.
//   ids    uint32[R]                      `point_list`
//   masks  uint32[slots * 8]              one word per (32-surfel batch of a tile, warp of the tile's CTA): bit j =
//                                         "some live pixel of the warp's 8x4 block may blend surfel j of the batch"
// Batch b of tile t (whose list starts at position x_t) owns slot  (x_t >> 5) + t + b : distinct for distinct (t, b)
// because floor(x/32) + ceil(n/32) <= floor((x+n)/32) + 1, and below R/32 + tiles + 1.
struct SbPairListLayout { size_t ids, masks, total; };
static inline __host__ SbPairListLayout sb_pair_list_layout(int64_t R, int W, int H) {
    SbPairListLayout L;
    size_t r = (size_t)(R > 0 ? R : 0);
    size_t tiles = (size_t)((W + SB_TILE_X - 1) / SB_TILE_X) * ((H + SB_TILE_Y - 1) / SB_TILE_Y);
    L.ids = 0;
    L.masks = sb_align(r * 4);
    L.total = sb_align(L.masks + (r / 32 + tiles + 2) * 8 * 4);
    return L;
}
static inline __host__ __device__ uint32_t sb_hit_slot(uint32_t list_begin, uint32_t tile) {
    return (list_begin >> 5) + tile;
}

// Binning scratch (caller-allocated, R-sized; only point_list survives the forward and it is a
// separate caller buffer).
struct SbBinLayout {
    size_t keys_a, keys_b;  // uint32[R]  tile id of every pair
    size_t vals_a, vals_b;  // uint32[R]  surfel id of every pair
    size_t chunk_first;     // uint32[ceil(R/1024)] depth-order index of the surfel owning each chunk's first pair
    size_t hist;            // uint32[SB_MAX_PASSES*256] exclusive digit bases of the pair sort (tile-count scan)
    size_t tickets;         // uint32[SB_MAX_PASSES]
    size_t status;          // uint32[passes * ntiles * 256] look-back words
    size_t zero_begin, zero_bytes; // tickets + status; a sort zeroes tickets + the status words it will use
    size_t total;
};

static inline __host__ SbBinLayout sb_bin_layout(int64_t R) {
    SbBinLayout L;
    size_t r = (size_t)(R > 0 ? R : 0);
    size_t ntiles = (r + SB_SORT_THREADS * SB_SORT_ITEMS_SMALL - 1) / (SB_SORT_THREADS * SB_SORT_ITEMS_SMALL);
    size_t o = 0;
    L.keys_a = o; o = sb_align(o + r * 4);
    L.keys_b = o; o = sb_align(o + r * 4);
    L.vals_a = o; o = sb_align(o + r * 4);
    L.vals_b = o; o = sb_align(o + r * 4);
    L.chunk_first = o; o = sb_align(o + ((r + SB_EMIT_CHUNK - 1) / SB_EMIT_CHUNK + 1) * 4);
    L.hist = o;    o = sb_align(o + SB_MAX_PASSES * SB_RADIX * 4);
    L.zero_begin = o;
    L.tickets = o; o = sb_align(o + SB_MAX_PASSES * 4);
    L.status = o;  o = sb_align(o + (size_t)SB_MAX_PASSES * ntiles * SB_RADIX * 4);
    L.zero_bytes = o - L.zero_begin;
    L.total = o;
    return L;
}

// Same bit count the reference passes to its sort: 32 + getHigherMsb(tiles)
// (DSR/cuda_rasterizer/rasterizer_impl.cu:35-50,410).
// bytes from BinLayout::zero_begin (or GeomLayout::tickets ...) a sort of n pairs in `passes` passes must find zeroed
static inline __host__ size_t sb_sort_status_bytes(int64_t n, int passes, int small) {
    size_t tile = (size_t)SB_SORT_THREADS * (small ? SB_SORT_ITEMS_SMALL : SB_SORT_ITEMS);
    return (size_t)passes * (((size_t)(n > 0 ? n : 0) + tile - 1) / tile) * SB_RADIX * 4;
}

static inline __host__ uint32_t sb_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// Masked Adam without bias correction, one scalar (DSR/cuda_rasterizer/adam.cu:26-37).  One definition for the three
// kernels that apply it (sb_adam_kernel, sb_surfel_update_kernel, sb_dp_post_kernel): the same expression tree, so
// nvcc contracts it the same way and the single-GPU pass, the NCCL path and the peer-memory path round identically.
__device__ __forceinline__ void sb_adam_scalar(float& p, float& m, float& v, float g, float lr, float b1, float b2, float eps) {
    m = b1 * m + (1.0f - b1) * g;
    v = b2 * v + (1.0f - b2) * g * g;
    p += -lr * m / (sqrtf(v) + eps);
}

#define SB_MAX_DEVICES 64
static inline int sb_current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= SB_MAX_DEVICES) dev = 0;
    return dev;
}

void sb_set_error(const char* fmt, ...);
void sb_count_launch(int n);  // kernels launched (reported as gpu_launches by bench.py)
// Optional per-stage device timing (api.cu): brackets a stage with CUDA events on its launching stream.
struct SbStageScope {
    bool on; int stage; cudaStream_t s; cudaEvent_t e0;
    SbStageScope(int stage, cudaStream_t s);
    ~SbStageScope();
};
#define SB_CUDA_CHECK(expr)                                                              \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            sb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return 2;                                                                    \
        }                                                                                \
    } while (0)

// Stage launchers (defined in the .cu files, sequenced by api.cu).
int sb_launch_preprocess(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                         const float* transmat_precomp, const float* opacities, const float* colors,
                         const float* viewmatrix, const float* projmatrix, char* geom, int* radii, cudaStream_t s);
int sb_launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s);
int sb_launch_surfel_sort(int P, char* geom, cudaStream_t s);
int sb_launch_duplicate(int P, int W, int H, int64_t R, int R_is_capacity, const char* geom, uint32_t* chunk_first,
                        uint32_t* tile_counts, uint32_t* keys, uint32_t* vals, cudaStream_t s);
int sb_launch_sort_hist(const uint32_t* keys, int64_t n, const SbSortPlan& pl, int drop_invalid, uint32_t* digit_base,
                        uint32_t* n_valid, cudaStream_t s);
int sb_launch_sort_passes(const SbSortPlan& pl, int small, const uint32_t* keys_in, const uint32_t* vals_in,
                          uint32_t* const kbuf[2], uint32_t* const vbuf[2], uint32_t* vals_final, int64_t n,
                          const uint32_t* n_dev, const uint32_t* digit_base, uint32_t* tickets, uint32_t* status,
                          cudaStream_t s);
int sb_launch_tile_scan(int tiles, const SbSortPlan& pl, const uint32_t* tile_counts, uint2* ranges, uint32_t* order,
                        uint32_t* digit_base, cudaStream_t s);
int sb_launch_blend_forward(int W, int H, const float* bg, const char* geom, const uint32_t* point_list,
                            char* image, float* out_color, float* out_others, uint32_t* hit_masks, cudaStream_t s);
int sb_launch_blend_backward(int W, int H, const float* bg, const char* geom, const uint32_t* point_list,
                             const char* image, const float* out_color, const float* out_others,
                             const float* dL_dcolor, const float* dL_dothers, const bool* pixel_mask,
                             float* grad_acc, const uint32_t* hit_masks, const uint32_t* walk_cap, cudaStream_t s);
int sb_launch_preprocess_backward(int P, int W, int H, float tanfovx, float tanfovy, const float* means3D,
                                  const float* scales, const float* rotations, const float* transmat_precomp,
                                  const float* viewmatrix,
                                  const float* projmatrix, const int* radii, const char* geom, const float* grad_acc,
                                  float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                                  float* dL_dscales, float* dL_drotations, float* dL_dtransMat, float* scores,
                                  cudaStream_t s);
size_t sb_map_loss_bytes(int W, int H);
int sb_launch_map_loss(int W, int H, const float* color, const float* allmap, const float* gt_rgb,
                       const float* gt_depth, const float* gt_cov, const float* w2c, float cx, float cy, float fx,
                       float fy, float w_rgb, float w_norm, float w_alpha, float w_depth, float w_dist, char* ws,
                       float* loss, float* dL_dcolor, float* dL_dallmap, cudaStream_t s);
int sb_launch_activate(int P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                       float* opacities, float* scales, float* rotations, cudaStream_t s);
int sb_launch_adam(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const bool* visible,
                   float lr, float b1, float b2, float eps, uint32_t N, uint32_t M, cudaStream_t s);
