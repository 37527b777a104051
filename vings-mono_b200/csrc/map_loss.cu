// N2: the mapping step's loss block, fused.
//
// Replaces, for one view, everything between the rasterizer's forward and backward in VINGS-Mono's
// train_once_gaussian (scripts/gaussian/gaussian_base.py:205-221 render_raw post-processing,
// scripts/gaussian/loss_utils.py:110-153 get_loss, :36-55 _ssim, scripts/gaussian/normal_utils.py:6-21,
// scripts/gaussian/tf.py:133-146) and the autograd backward of all of it: ~170 torch kernels (5 depthwise
// 11x11 convolutions and their backward, boolean-mask gathers whose backward is a sort + scatter, ~30
// image-sized elementwise passes) become 5 launches that read the rendered maps and the ground truth
// once or twice and write dL/dcolor[3,H,W] and dL/dallmap[7,H,W] directly.
//
//   total = w_rgb  * (0.8 * mean_valid|rgb - gt| + 0.2 * (1 - mean_valid SSIM(rgb, gt)))
//         + w_norm * mean_valid(1 - <n_world, surf_n_world>)
//         + w_alpha* mean_sky(alpha)
//         + w_depth* mean_valid(|depth - gt_depth| / depth_cov)
//         + w_dist * mean_valid(distortion)
//   sky   = (sum_c gt_rgb == 0),  valid = !sky && gt_depth > 0                      (loss_utils.py:112-113)
//   depth = nan_to_num(allmap[0] / allmap[1], 0, 0)                                 (gaussian_base.py:209-210)
//   n_world = allmap[2:5] @ w2c[:3,:3],  surf_n = normalize(cross(l2r, b2t)) @ w2c[:3,:3], 0 on the border
//
// Kernels (all HBM-streaming; the whole block moves ~60 B/pixel/launch):
//   sb_loss_count_kernel     pixel counts N_valid, N_sky (the means' denominators)
//   sb_ssim_stats_kernel     separable 11x11 Gaussian statistics in shared memory (32x16 tile + halo 5),
//                            SSIM value, L1 value, and the three per-pixel derivative maps
//                            d ssim / d mu1, d ssim / d E[x^2], d ssim / d E[xy]
//   sb_ssim_grad_kernel      the same separable filter over the three maps -> dL/dcolor (+ the L1 term)
//   sb_geom_loss_kernel      depth / alpha / normal / distortion terms and dL/dallmap, including the
//                            backward of the depth -> normal stencil as a gather (no atomics)
//   sb_loss_final_kernel     the scalar
#include "common.cuh"
#include <mutex>

#define ML_TX 32
#define ML_TY 16
#define ML_HALO 5
#define ML_RX (ML_TX + 2 * ML_HALO)  // 42
#define ML_RY (ML_TY + 2 * ML_HALO)  // 26
#define ML_THREADS (ML_TX * ML_TY)   // 512

// scratch header (zeroed per call), followed by the 9 derivative maps [3 maps][3 channels][HW]
struct SbLossHeader {
    unsigned int n_valid, n_sky;
    double acc[6];  // 0 sum|rgb-gt|  1 sum ssim  2 sum (1 - n.sn)  3 sum alpha over sky  4 sum |d-gt|/cov  5 sum dist
};
#define ML_HEADER_BYTES 256

size_t sb_map_loss_bytes(int W, int H) {
    return sb_align(ML_HEADER_BYTES + (size_t)W * H * 9 * sizeof(float));
}

// 11-tap Gaussian, sigma 1.5, normalised (loss_utils.py:26-28).  Filled by the launcher in float the way
// the reference builds it: exp in double -> float tensor -> divided by its float sum.
__constant__ float c_gauss[11];

__device__ __forceinline__ bool ml_sky(const float* gt_rgb, size_t HW, size_t p) {
    return (gt_rgb[p] + gt_rgb[p + HW] + gt_rgb[p + 2 * HW]) == 0.0f;
}

__global__ void sb_loss_count_kernel(int HW, const float* __restrict__ gt_rgb, const float* __restrict__ gt_depth,
                                     SbLossHeader* __restrict__ hdr) {
    unsigned int nv = 0, ns = 0;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const bool sky = ml_sky(gt_rgb, (size_t)HW, (size_t)p);
        ns += sky;
        nv += (!sky && gt_depth[p] > 0.0f);
    }
    nv = __reduce_add_sync(0xffffffffu, nv);
    ns = __reduce_add_sync(0xffffffffu, ns);
    if ((threadIdx.x & 31) == 0) {
        if (nv) atomicAdd(&hdr->n_valid, nv);
        if (ns) atomicAdd(&hdr->n_sky, ns);
    }
}

// block-wide sum of a double, result valid in thread 0
template <int NWARPS>
__device__ __forceinline__ double ml_block_sum(double v, double* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 0; w < NWARPS; w++) t += s_red[w];
    }
    return t;
}

// ---- SSIM statistics + derivative maps (loss_utils.py:36-55) ----
// grid (tiles_x, tiles_y, 3 channels), 512 threads: thread (tx, ty) owns pixel (x0+tx, y0+ty).
__global__ void __launch_bounds__(ML_THREADS)
sb_ssim_stats_kernel(int W, int H, const float* __restrict__ img1, const float* __restrict__ img2,
                     const float* __restrict__ gt_depth, SbLossHeader* __restrict__ hdr, float* __restrict__ maps) {
    __shared__ float s_x[ML_RY][ML_RX + 1];
    __shared__ float s_y[ML_RY][ML_RX + 1];
    __shared__ float s_h[5][ML_RY][ML_TX + 1];
    __shared__ double s_red[ML_THREADS / 32];
    const int tid = threadIdx.x, tx = tid & (ML_TX - 1), ty = tid / ML_TX;
    const int x0 = blockIdx.x * ML_TX, y0 = blockIdx.y * ML_TY, ch = blockIdx.z;
    const size_t HW = (size_t)W * H;
    const float* X = img1 + ch * HW;
    const float* Y = img2 + ch * HW;
    // zero padding outside the image, like conv2d(padding=5)
    for (int e = tid; e < ML_RY * ML_RX; e += ML_THREADS) {
        const int ry = e / ML_RX, rx = e - ry * ML_RX;
        const int gx = x0 + rx - ML_HALO, gy = y0 + ry - ML_HALO;
        float a = 0.f, b = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            a = X[(size_t)gy * W + gx];
            b = Y[(size_t)gy * W + gx];
        }
        s_x[ry][rx] = a;
        s_y[ry][rx] = b;
    }
    __syncthreads();
    // horizontal pass: rows of the region x columns of the tile
    for (int e = tid; e < ML_RY * ML_TX; e += ML_THREADS) {
        const int ry = e / ML_TX, cx = e - ry * ML_TX;
        float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_gauss[k], a = s_x[ry][cx + k], b = s_y[ry][cx + k];
            m1 += w * a;
            m2 += w * b;
            e11 += w * a * a;
            e22 += w * b * b;
            e12 += w * a * b;
        }
        s_h[0][ry][cx] = m1; s_h[1][ry][cx] = m2; s_h[2][ry][cx] = e11; s_h[3][ry][cx] = e22; s_h[4][ry][cx] = e12;
    }
    __syncthreads();
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float w = c_gauss[k];
        mu1 += w * s_h[0][ty + k][tx];
        mu2 += w * s_h[1][ty + k][tx];
        e11 += w * s_h[2][ty + k][tx];
        e22 += w * s_h[3][ty + k][tx];
        e12 += w * s_h[4][ty + k][tx];
    }
    const int gx = x0 + tx, gy = y0 + ty;
    double l1 = 0.0, ss = 0.0;
    if (gx < W && gy < H) {
        const size_t p = (size_t)gy * W + gx;
        const bool valid = !ml_sky(img2, HW, p) && gt_depth[p] > 0.0f;
        float d_mu1 = 0.f, d_e11 = 0.f, d_e12 = 0.f;
        if (valid) {
            const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float sig1 = e11 - mu1_sq, sig2 = e22 - mu2_sq, sig12 = e12 - mu12;
            const float A = mu1_sq + mu2_sq + C1, B = sig1 + sig2 + C2;
            const float C = 2.f * mu12 + C1, D = 2.f * sig12 + C2;
            const float iAB = 1.0f / (A * B);
            ss = (double)(C * D * iAB);
            // total derivative w.r.t. mu1 (through A, B, C, D), and partials w.r.t. E[x^2], E[xy]
            d_mu1 = 2.f * iAB * (mu2 * (D - C) - mu1 * C * D * (1.0f / A - 1.0f / B));
            d_e11 = -C * D * iAB / B;
            d_e12 = 2.f * C * iAB;
            l1 = (double)fabsf(s_x[ty + ML_HALO][tx + ML_HALO] - s_y[ty + ML_HALO][tx + ML_HALO]);
        }
        maps[(0 * 3 + ch) * HW + p] = d_mu1;
        maps[(1 * 3 + ch) * HW + p] = d_e11;
        maps[(2 * 3 + ch) * HW + p] = d_e12;
    }
    const double t1 = ml_block_sum<ML_THREADS / 32>(l1, s_red);
    const double t2 = ml_block_sum<ML_THREADS / 32>(ss, s_red);
    if (tid == 0) {
        if (t1 != 0.0) atomicAdd(&hdr->acc[0], t1);
        if (t2 != 0.0) atomicAdd(&hdr->acc[1], t2);
    }
}

// ---- dL/dcolor: filter the derivative maps back (the adjoint of a symmetric zero-padded filter is itself) ----
__global__ void __launch_bounds__(ML_THREADS)
sb_ssim_grad_kernel(int W, int H, const float* __restrict__ img1, const float* __restrict__ img2,
                    const float* __restrict__ gt_depth, const SbLossHeader* __restrict__ hdr,
                    const float* __restrict__ maps, float w_rgb, float* __restrict__ dL_dcolor) {
    __shared__ float s_m[3][ML_RY][ML_RX + 1];
    __shared__ float s_h[3][ML_RY][ML_TX + 1];
    const int tid = threadIdx.x, tx = tid & (ML_TX - 1), ty = tid / ML_TX;
    const int x0 = blockIdx.x * ML_TX, y0 = blockIdx.y * ML_TY, ch = blockIdx.z;
    const size_t HW = (size_t)W * H;
    for (int e = tid; e < ML_RY * ML_RX; e += ML_THREADS) {
        const int ry = e / ML_RX, rx = e - ry * ML_RX;
        const int gx = x0 + rx - ML_HALO, gy = y0 + ry - ML_HALO;
        float a = 0.f, b = 0.f, c = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t p = (size_t)gy * W + gx;
            a = maps[(0 * 3 + ch) * HW + p];
            b = maps[(1 * 3 + ch) * HW + p];
            c = maps[(2 * 3 + ch) * HW + p];
        }
        s_m[0][ry][rx] = a; s_m[1][ry][rx] = b; s_m[2][ry][rx] = c;
    }
    __syncthreads();
    for (int e = tid; e < ML_RY * ML_TX; e += ML_THREADS) {
        const int ry = e / ML_TX, cx = e - ry * ML_TX;
        float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_gauss[k];
            a += w * s_m[0][ry][cx + k];
            b += w * s_m[1][ry][cx + k];
            c += w * s_m[2][ry][cx + k];
        }
        s_h[0][ry][cx] = a; s_h[1][ry][cx] = b; s_h[2][ry][cx] = c;
    }
    __syncthreads();
    const int gx = x0 + tx, gy = y0 + ty;
    if (gx >= W || gy >= H) return;
    float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float w = c_gauss[k];
        a += w * s_h[0][ty + k][tx];
        b += w * s_h[1][ty + k][tx];
        c += w * s_h[2][ty + k][tx];
    }
    const size_t p = (size_t)gy * W + gx;
    const float x = img1[ch * HW + p], y = img2[ch * HW + p];
    const unsigned int nv = hdr->n_valid;
    float g = 0.f;
    if (nv > 0) {
        const float inv = w_rgb / (3.0f * (float)nv);
        g = -0.2f * inv * (a + 2.f * x * b + y * c);
        const bool valid = !ml_sky(img2, HW, p) && gt_depth[p] > 0.0f;
        if (valid) {
            const float d = x - y;
            g += 0.8f * inv * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
        }
    }
    dL_dcolor[ch * HW + p] = g;
}

// ---- depth / alpha / normal / distortion terms and dL/dallmap ----
#define GL_TX 32
#define GL_TY 8
#define GL_THREADS (GL_TX * GL_TY)

struct SbLossCam {
    float cx, cy, fx, fy;  // camera point of pixel (u, v) at depth d: ((u-cx)/fx*d, (v-cy)/fy*d, d)
    float w_norm, w_alpha, w_depth, w_dist;
};

// depth = nan_to_num(a0 / a1, nan=0, posinf=0) (gaussian_base.py:209-210)
__device__ __forceinline__ float ml_depth(float a0, float a1, bool* ok) {
    const float d = a0 / a1;
    const bool fin = isfinite(d);
    *ok = fin;
    return fin ? d : 0.0f;
}

template <bool NORMAL>
__global__ void __launch_bounds__(GL_THREADS)
sb_geom_loss_kernel(int W, int H, const float* __restrict__ allmap, const float* __restrict__ gt_rgb,
                    const float* __restrict__ gt_depth, const float* __restrict__ gt_cov,
                    const float* __restrict__ w2c, SbLossCam cam, SbLossHeader* __restrict__ hdr,
                    float* __restrict__ dL_dallmap) {
    __shared__ float s_d[GL_TY + 4][GL_TX + 4];          // depth, halo 2
    __shared__ float s_c[9][GL_TY + 2][GL_TX + 2];       // per centre (halo 1): da[3], db[3], n[3]
    __shared__ double s_red[GL_THREADS / 32];
    const int tid = threadIdx.x, tx = tid & (GL_TX - 1), ty = tid / GL_TX;
    const int x0 = blockIdx.x * GL_TX, y0 = blockIdx.y * GL_TY;
    const size_t HW = (size_t)W * H;
    const unsigned int nv = hdr->n_valid, ns = hdr->n_sky;
    const float inv_nv = nv ? 1.0f / (float)nv : 0.0f;
    const float inv_ns = ns ? 1.0f / (float)ns : 0.0f;
    float Rm[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) Rm[i][j] = NORMAL ? w2c[i * 4 + j] : 0.f;

    if (NORMAL) {
        for (int e = tid; e < (GL_TY + 4) * (GL_TX + 4); e += GL_THREADS) {
            const int ry = e / (GL_TX + 4), rx = e - ry * (GL_TX + 4);
            const int gx = x0 + rx - 2, gy = y0 + ry - 2;
            float d = 0.f;
            if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
                const size_t p = (size_t)gy * W + gx;
                bool ok;
                d = ml_depth(allmap[p], allmap[HW + p], &ok);
            }
            s_d[ry][rx] = d;
        }
        __syncthreads();
        for (int e = tid; e < (GL_TY + 2) * (GL_TX + 2); e += GL_THREADS) {
            const int ry = e / (GL_TX + 2), rx = e - ry * (GL_TX + 2);
            const int u = x0 + rx - 1, v = y0 + ry - 1;
            float da[3] = {0.f, 0.f, 0.f}, db[3] = {0.f, 0.f, 0.f}, n[3] = {0.f, 0.f, 0.f};
            if (u >= 1 && u <= W - 2 && v >= 1 && v <= H - 2) {  // normal_utils.py:12-15: interior only
                const float Xl = ((float)(u - 1) - cam.cx) / cam.fx, Xr = ((float)(u + 1) - cam.cx) / cam.fx;
                const float Xc = ((float)u - cam.cx) / cam.fx;
                const float Yt = ((float)(v - 1) - cam.cy) / cam.fy, Yb = ((float)(v + 1) - cam.cy) / cam.fy;
                const float Yc = ((float)v - cam.cy) / cam.fy;
                const float dl = s_d[ry + 1][rx], dr = s_d[ry + 1][rx + 2], dt = s_d[ry][rx + 1], dbm = s_d[ry + 2][rx + 1];
                const float a[3] = {Xr * dr - Xl * dl, Yc * dr - Yc * dl, dr - dl};      // left -> right
                const float b[3] = {Xc * dt - Xc * dbm, Yt * dt - Yb * dbm, dt - dbm};   // bottom -> top
                const float c[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
                const float len = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
                const float den = fmaxf(len, 1e-12f);
                n[0] = c[0] / den; n[1] = c[1] / den; n[2] = c[2] / den;
                const size_t p = (size_t)v * W + u;
                const bool valid = !ml_sky(gt_rgb, HW, p) && gt_depth[p] > 0.0f;
                if (valid && cam.w_norm != 0.f) {
                    const float an[3] = {allmap[2 * HW + p], allmap[3 * HW + p], allmap[4 * HW + p]};
                    float rw[3];
#pragma unroll
                    for (int j = 0; j < 3; j++) rw[j] = an[0] * Rm[0][j] + an[1] * Rm[1][j] + an[2] * Rm[2][j];
                    const float g = cam.w_norm * inv_nv;
                    float dn[3];
#pragma unroll
                    for (int i = 0; i < 3; i++) dn[i] = -g * (Rm[i][0] * rw[0] + Rm[i][1] * rw[1] + Rm[i][2] * rw[2]);
                    float dc[3];
                    if (len > 1e-12f) {
                        const float nd = n[0] * dn[0] + n[1] * dn[1] + n[2] * dn[2];
#pragma unroll
                        for (int i = 0; i < 3; i++) dc[i] = (dn[i] - n[i] * nd) / len;
                    } else {
#pragma unroll
                        for (int i = 0; i < 3; i++) dc[i] = dn[i] / 1e-12f;
                    }
                    // c = a x b  ->  da = b x dc, db = dc x a
                    da[0] = b[1] * dc[2] - b[2] * dc[1]; da[1] = b[2] * dc[0] - b[0] * dc[2]; da[2] = b[0] * dc[1] - b[1] * dc[0];
                    db[0] = dc[1] * a[2] - dc[2] * a[1]; db[1] = dc[2] * a[0] - dc[0] * a[2]; db[2] = dc[0] * a[1] - dc[1] * a[0];
                }
            }
#pragma unroll
            for (int i = 0; i < 3; i++) {
                s_c[i][ry][rx] = da[i];
                s_c[3 + i][ry][rx] = db[i];
                s_c[6 + i][ry][rx] = n[i];
            }
        }
        __syncthreads();
    }

    const int u = x0 + tx, v = y0 + ty;
    double t_norm = 0.0, t_alpha = 0.0, t_depth = 0.0, t_dist = 0.0;
    if (u < W && v < H) {
        const size_t p = (size_t)v * W + u;
        const bool sky = ml_sky(gt_rgb, HW, p);
        const float gtd = gt_depth[p];
        const bool valid = !sky && gtd > 0.0f;
        const float a0 = allmap[p], a1 = allmap[HW + p];
        bool ok;
        const float d = ml_depth(a0, a1, &ok);
        float g_d = 0.f;                      // dL/d depth
        float g_n[3] = {0.f, 0.f, 0.f};       // dL/d allmap[2:5]
        float g_a1 = sky ? cam.w_alpha * inv_ns : 0.f;
        if (sky) t_alpha = (double)a1;
        if (valid) {
            const float w = 1.0f / gt_cov[p];
            const float diff = d - gtd;
            t_depth = (double)(fabsf(diff) * w);
            g_d = cam.w_depth * inv_nv * w * (diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f));
            t_dist = (double)allmap[6 * HW + p];
        }
        if (NORMAL) {
            const int cy_ = ty + 1, cx_ = tx + 1;
            if (valid) {
                const float n[3] = {s_c[6][cy_][cx_], s_c[7][cy_][cx_], s_c[8][cy_][cx_]};
                const float an[3] = {allmap[2 * HW + p], allmap[3 * HW + p], allmap[4 * HW + p]};
                float rw[3], sw[3];
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    rw[j] = an[0] * Rm[0][j] + an[1] * Rm[1][j] + an[2] * Rm[2][j];
                    sw[j] = n[0] * Rm[0][j] + n[1] * Rm[1][j] + n[2] * Rm[2][j];
                }
                t_norm = (double)(1.0f - (rw[0] * sw[0] + rw[1] * sw[1] + rw[2] * sw[2]));
                const float g = cam.w_norm * inv_nv;
#pragma unroll
                for (int i = 0; i < 3; i++) g_n[i] = -g * (Rm[i][0] * sw[0] + Rm[i][1] * sw[1] + Rm[i][2] * sw[2]);
            }
            // gather the stencil's adjoint: this pixel is the right / left / top / bottom neighbour of 4 centres
            float dpc[3];
#pragma unroll
            for (int i = 0; i < 3; i++)
                dpc[i] = s_c[i][cy_][cx_ - 1] - s_c[i][cy_][cx_ + 1] + s_c[3 + i][cy_ + 1][cx_] - s_c[3 + i][cy_ - 1][cx_];
            const float Xc = ((float)u - cam.cx) / cam.fx, Yc = ((float)v - cam.cy) / cam.fy;
            g_d += dpc[0] * Xc + dpc[1] * Yc + dpc[2];
        }
        // depth = a0 / a1 (where finite): torch yields NaN at a1 == 0; nothing is blended there, 0 is written
        float g_a0 = 0.f;
        if (ok) {
            g_a0 = g_d / a1;
            g_a1 += -g_d * d / a1;
        }
        dL_dallmap[p] = g_a0;
        dL_dallmap[HW + p] = g_a1;
        dL_dallmap[2 * HW + p] = g_n[0];
        dL_dallmap[3 * HW + p] = g_n[1];
        dL_dallmap[4 * HW + p] = g_n[2];
        dL_dallmap[5 * HW + p] = 0.f;   // median depth: unused by the loss
        dL_dallmap[6 * HW + p] = valid ? cam.w_dist * inv_nv : 0.f;
    }
    const double r0 = ml_block_sum<GL_THREADS / 32>(t_norm, s_red);
    const double r1 = ml_block_sum<GL_THREADS / 32>(t_alpha, s_red);
    const double r2 = ml_block_sum<GL_THREADS / 32>(t_depth, s_red);
    const double r3 = ml_block_sum<GL_THREADS / 32>(t_dist, s_red);
    if (tid == 0) {
        if (r0 != 0.0) atomicAdd(&hdr->acc[2], r0);
        if (r1 != 0.0) atomicAdd(&hdr->acc[3], r1);
        if (r2 != 0.0) atomicAdd(&hdr->acc[4], r2);
        if (r3 != 0.0) atomicAdd(&hdr->acc[5], r3);
    }
}

__global__ void sb_loss_final_kernel(const SbLossHeader* __restrict__ hdr, float w_rgb, float w_norm, float w_alpha,
                                     float w_depth, float w_dist, float* __restrict__ loss) {
    // means over empty sets are NaN, as torch's .mean() of an empty selection is (loss_utils.py:13,130)
    const float nv = (float)hdr->n_valid, ns = (float)hdr->n_sky;
    const float l1 = (float)hdr->acc[0] / (3.0f * nv);
    const float ssim = (float)hdr->acc[1] / (3.0f * nv);
    const float rgb_loss = 0.8f * l1 + 0.2f * (1.0f - ssim);
    const float normal_loss = (float)hdr->acc[2] / nv;
    const float alpha_loss = (float)hdr->acc[3] / ns;
    const float depth_loss = (float)hdr->acc[4] / nv;
    const float dist_loss = (float)hdr->acc[5] / nv;
    loss[0] = w_rgb * rgb_loss + w_norm * normal_loss + w_alpha * alpha_loss + w_depth * depth_loss + w_dist * dist_loss;
    // the six means, for logging / tests
    loss[1] = l1; loss[2] = ssim; loss[3] = normal_loss; loss[4] = alpha_loss; loss[5] = depth_loss; loss[6] = dist_loss;
}

int sb_launch_map_loss(int W, int H, const float* color, const float* allmap, const float* gt_rgb,
                       const float* gt_depth, const float* gt_cov, const float* w2c, float cx, float cy, float fx,
                       float fy, float w_rgb, float w_norm, float w_alpha, float w_depth, float w_dist, char* ws,
                       float* loss, float* dL_dcolor, float* dL_dallmap, cudaStream_t s) {
    // same constants for every call; constant memory is per device, so the upload is tracked per device
    // (cudaMemcpyToSymbol writes the CURRENT device's copy only) and guarded for concurrent first calls
    static std::mutex gauss_mu;
    static bool gauss_ready_dev[SB_MAX_DEVICES] = {};
    const int gdev = sb_current_device();
    std::lock_guard<std::mutex> gauss_lock(gauss_mu);
    if (!gauss_ready_dev[gdev]) {
        float g[11], sum = 0.f;
        for (int x = 0; x < 11; x++) {
            g[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5));
            sum += g[x];
        }
        for (int x = 0; x < 11; x++) g[x] /= sum;
        SB_CUDA_CHECK(cudaMemcpyToSymbol(c_gauss, g, sizeof(g)));
        gauss_ready_dev[gdev] = true;
    }
    const int HW = W * H;
    SbLossHeader* hdr = (SbLossHeader*)ws;
    float* maps = (float*)(ws + ML_HEADER_BYTES);
    SB_CUDA_CHECK(cudaMemsetAsync(hdr, 0, ML_HEADER_BYTES, s));
    sb_loss_count_kernel<<<148 * 4, 256, 0, s>>>(HW, gt_rgb, gt_depth, hdr);
    const dim3 sg((W + ML_TX - 1) / ML_TX, (H + ML_TY - 1) / ML_TY, 3);
    sb_ssim_stats_kernel<<<sg, ML_THREADS, 0, s>>>(W, H, color, gt_rgb, gt_depth, hdr, maps);
    sb_ssim_grad_kernel<<<sg, ML_THREADS, 0, s>>>(W, H, color, gt_rgb, gt_depth, hdr, maps, w_rgb, dL_dcolor);
    const dim3 gg((W + GL_TX - 1) / GL_TX, (H + GL_TY - 1) / GL_TY);
    SbLossCam cam = {cx, cy, fx, fy, w_norm, w_alpha, w_depth, w_dist};
    if (w_norm != 0.f)
        sb_geom_loss_kernel<true><<<gg, GL_THREADS, 0, s>>>(W, H, allmap, gt_rgb, gt_depth, gt_cov, w2c, cam, hdr, dL_dallmap);
    else
        sb_geom_loss_kernel<false><<<gg, GL_THREADS, 0, s>>>(W, H, allmap, gt_rgb, gt_depth, gt_cov, w2c, cam, hdr, dL_dallmap);
    sb_loss_final_kernel<<<1, 1, 0, s>>>(hdr, w_rgb, w_norm, w_alpha, w_depth, w_dist, loss);
    sb_count_launch(5);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
