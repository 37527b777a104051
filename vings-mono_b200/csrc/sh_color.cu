// View-dependent colour from spherical harmonics (degree 0..3) and its backward: the `shs` / `sh_degree` inputs of
// the rasterizer that VINGS-Mono itself never uses (it always passes colors_precomp).
//
// Replaces computeColorFromSH of DSR/cuda_rasterizer/forward.cu:28-79 (called from preprocessCUDA, :237-242) and
// backward.cu:20-139 (called from preprocessCUDA, :1038-1039).  Here the conversion is a stage of its own in front
// of / behind the rasterizer (colour = sum_k basis_k(dir) * sh_k + 0.5, clamped at 0), so the hot preprocess kernels
// stay free of it: sb200_sh_forward writes colors[P,3] (what the rasterizer takes as colors_precomp) and the clamp
// flags, sb200_sh_backward turns the rasterizer's dL/dcolors into dL/dsh and adds the view-direction term to
// dL/dmeans3D.  The real SH basis and its gradient are written out once (sh_basis) and used by both directions.
#include "common.cuh"
#include "../../include/surfel_b200.h"

#define SH_MAX_COEFFS 16

// basis values b[0..n) of the real spherical harmonics in the sign convention of the 3DGS family, and (GRAD) their
// partial derivatives with respect to the (unit) direction
template <bool GRAD>
__device__ __forceinline__ void sh_basis(int degree, float x, float y, float z, float* b, float3* g) {
    b[0] = 0.28209479177387814f;
    if (GRAD) g[0] = make_float3(0.f, 0.f, 0.f);
    if (degree < 1) return;
    const float c1 = 0.4886025119029199f;
    b[1] = -c1 * y; b[2] = c1 * z; b[3] = -c1 * x;
    if (GRAD) { g[1] = make_float3(0.f, -c1, 0.f); g[2] = make_float3(0.f, 0.f, c1); g[3] = make_float3(-c1, 0.f, 0.f); }
    if (degree < 2) return;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    const float c20 = 1.0925484305920792f, c22 = 0.31539156525252005f, c24 = 0.5462742152960396f;
    b[4] = c20 * xy; b[5] = -c20 * yz; b[6] = c22 * (2.0f * zz - xx - yy); b[7] = -c20 * xz; b[8] = c24 * (xx - yy);
    if (GRAD) {
        g[4] = make_float3(c20 * y, c20 * x, 0.f);
        g[5] = make_float3(0.f, -c20 * z, -c20 * y);
        g[6] = make_float3(c22 * -2.0f * x, c22 * -2.0f * y, c22 * 4.0f * z);
        g[7] = make_float3(-c20 * z, 0.f, -c20 * x);
        g[8] = make_float3(c24 * 2.0f * x, c24 * -2.0f * y, 0.f);
    }
    if (degree < 3) return;
    const float c30 = -0.5900435899266435f, c31 = 2.890611442640554f, c32 = -0.4570457994644658f,
                c33 = 0.3731763325901154f, c35 = 1.445305721320277f;
    b[9] = c30 * y * (3.0f * xx - yy);
    b[10] = c31 * xy * z;
    b[11] = c32 * y * (4.0f * zz - xx - yy);
    b[12] = c33 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
    b[13] = c32 * x * (4.0f * zz - xx - yy);
    b[14] = c35 * z * (xx - yy);
    b[15] = c30 * x * (xx - 3.0f * yy);
    if (GRAD) {
        g[9] = make_float3(c30 * 6.0f * xy, c30 * 3.0f * (xx - yy), 0.f);
        g[10] = make_float3(c31 * yz, c31 * xz, c31 * xy);
        g[11] = make_float3(c32 * -2.0f * xy, c32 * (4.0f * zz - xx - 3.0f * yy), c32 * 8.0f * yz);
        g[12] = make_float3(c33 * -6.0f * xz, c33 * -6.0f * yz, c33 * 3.0f * (2.0f * zz - xx - yy));
        g[13] = make_float3(c32 * (4.0f * zz - 3.0f * xx - yy), c32 * -2.0f * xy, c32 * 8.0f * xz);
        g[14] = make_float3(c35 * 2.0f * xz, c35 * -2.0f * yz, c35 * (xx - yy));
        g[15] = make_float3(c30 * 3.0f * (xx - yy), c30 * -6.0f * xy, 0.f);
    }
}

__global__ void __launch_bounds__(256)
sb_sh_forward_kernel(int P, int degree, int n_coeffs, const float* __restrict__ means3D, const float* __restrict__ campos,
                     const float* __restrict__ shs, float* __restrict__ colors, uint8_t* __restrict__ clamped) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float dx = means3D[3 * (size_t)i] - campos[0], dy = means3D[3 * (size_t)i + 1] - campos[1],
                dz = means3D[3 * (size_t)i + 2] - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float b[SH_MAX_COEFFS];
    sh_basis<false>(degree, dx / len, dy / len, dz / len, b, nullptr);
    const int n = (degree + 1) * (degree + 1);
    const float* sh = shs + (size_t)i * n_coeffs * 3;
    float r = 0.f, g = 0.f, bl = 0.f;
#pragma unroll
    for (int k = 0; k < SH_MAX_COEFFS; k++) {
        if (k < n) {
            r += b[k] * sh[3 * k];
            g += b[k] * sh[3 * k + 1];
            bl += b[k] * sh[3 * k + 2];
        }
    }
    r += 0.5f; g += 0.5f; bl += 0.5f;
    // colours are clamped to >= 0; the backward needs to know where (the clamp has zero gradient)
    clamped[3 * (size_t)i] = r < 0.f;
    clamped[3 * (size_t)i + 1] = g < 0.f;
    clamped[3 * (size_t)i + 2] = bl < 0.f;
    colors[3 * (size_t)i] = fmaxf(r, 0.f);
    colors[3 * (size_t)i + 1] = fmaxf(g, 0.f);
    colors[3 * (size_t)i + 2] = fmaxf(bl, 0.f);
}

__global__ void __launch_bounds__(256)
sb_sh_backward_kernel(int P, int degree, int n_coeffs, const float* __restrict__ means3D,
                      const float* __restrict__ campos, const float* __restrict__ shs,
                      const uint8_t* __restrict__ clamped, const int* __restrict__ radii,
                      const float* __restrict__ dL_dcolors, float* __restrict__ dL_dmeans3D,
                      float* __restrict__ dL_dsh) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float* out = dL_dsh + (size_t)i * n_coeffs * 3;
    if (radii[i] <= 0) {  // the reference's backward preprocess skips surfels that touched no tile: zero rows
        for (int k = 0; k < n_coeffs * 3; k++) out[k] = 0.f;
        return;
    }
    const float dx = means3D[3 * (size_t)i] - campos[0], dy = means3D[3 * (size_t)i + 1] - campos[1],
                dz = means3D[3 * (size_t)i + 2] - campos[2];
    const float len2 = dx * dx + dy * dy + dz * dz, len = sqrtf(len2);
    float b[SH_MAX_COEFFS];
    float3 gb[SH_MAX_COEFFS];
    sh_basis<true>(degree, dx / len, dy / len, dz / len, b, gb);
    const int n = (degree + 1) * (degree + 1);
    const float* sh = shs + (size_t)i * n_coeffs * 3;
    const float dr = clamped[3 * (size_t)i] ? 0.f : dL_dcolors[3 * (size_t)i];
    const float dg = clamped[3 * (size_t)i + 1] ? 0.f : dL_dcolors[3 * (size_t)i + 1];
    const float db = clamped[3 * (size_t)i + 2] ? 0.f : dL_dcolors[3 * (size_t)i + 2];
    float3 ddir = make_float3(0.f, 0.f, 0.f);  // dL/d(unit direction)
#pragma unroll
    for (int k = 0; k < SH_MAX_COEFFS; k++) {
        if (k < n) {
            out[3 * k] = b[k] * dr;
            out[3 * k + 1] = b[k] * dg;
            out[3 * k + 2] = b[k] * db;
            const float w = sh[3 * k] * dr + sh[3 * k + 1] * dg + sh[3 * k + 2] * db;
            ddir.x += gb[k].x * w; ddir.y += gb[k].y * w; ddir.z += gb[k].z * w;
        }
    }
    for (int k = 3 * n; k < n_coeffs * 3; k++) out[k] = 0.f;  // coefficients above the active degree
    // through the normalisation dir = d / |d|:  (|d|^2 I - d d^T) ddir / |d|^3
    const float inv3 = 1.0f / (len2 * len);
    const float dot = dx * ddir.x + dy * ddir.y + dz * ddir.z;
    dL_dmeans3D[3 * (size_t)i] += (len2 * ddir.x - dx * dot) * inv3;
    dL_dmeans3D[3 * (size_t)i + 1] += (len2 * ddir.y - dy * dot) * inv3;
    dL_dmeans3D[3 * (size_t)i + 2] += (len2 * ddir.z - dz * dot) * inv3;
}

static int sh_check(int P, int degree, int n_coeffs) {
    if (P < 0 || degree < 0 || degree > 3 || n_coeffs < (degree + 1) * (degree + 1)) {
        sb_set_error("sh: bad P / degree %d / coefficient count %d", degree, n_coeffs);
        return 1;
    }
    return 0;
}

extern "C" {

int sb200_sh_forward(int P, int degree, int n_coeffs, const float* means3D, const float* campos, const float* shs,
                     float* colors, uint8_t* clamped, void* stream) {
    if (sh_check(P, degree, n_coeffs)) return 1;
    if (P == 0) return 0;
    if (!means3D || !campos || !shs || !colors || !clamped) { sb_set_error("sh_forward: NULL input"); return 1; }
    SbStageScope sc(SB200_STAGE_PREPROCESS, (cudaStream_t)stream);
    sb_sh_forward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, degree, n_coeffs, means3D, campos, shs,
                                                                           colors, clamped);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int sb200_sh_backward(int P, int degree, int n_coeffs, const float* means3D, const float* campos, const float* shs,
                      const uint8_t* clamped, const int* radii, const float* dL_dcolors, float* dL_dmeans3D,
                      float* dL_dsh, void* stream) {
    if (sh_check(P, degree, n_coeffs)) return 1;
    if (P == 0) return 0;
    if (!means3D || !campos || !shs || !clamped || !radii || !dL_dcolors || !dL_dmeans3D || !dL_dsh) {
        sb_set_error("sh_backward: NULL input");
        return 1;
    }
    SbStageScope sc(SB200_STAGE_PREPROCESS_BWD, (cudaStream_t)stream);
    sb_sh_backward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        P, degree, n_coeffs, means3D, campos, shs, clamped, radii, dL_dcolors, dL_dmeans3D, dL_dsh);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // extern "C"
