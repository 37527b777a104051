// N1: everything the mapping step does per surfel around the rasterizer, fused.
//
//   sb_activate_kernel        raw parameters -> rasterizer inputs: sigmoid(opacity), exp(scaling),
//                             normalize(rotation)                       (scripts/gaussian/gaussian_base.py:160-166)
//   sb_surfel_update_kernel   one pass over the surfels after the rasterizer's backward:
//       backward of those three activations (what autograd does for the reference),
//       add_records score bookkeeping              (scripts/gaussian/gaussian_model.py:163-168, gaussian_base.py:388-392)
//       weighting_grad                             (scripts/gaussian/gaussian_utils.py:56-62)
//       radii[_stable_mask] = 0; SparseGaussianAdam.step(radii > 0) for the five parameter groups
//                                                  (gaussian_base.py:400-403, DSR/cuda_rasterizer/adam.cu:9-38)
// The reference runs ~45 P-sized torch kernels plus five Adam launches for this; here every array is read
// and written once (13 parameters + 26 moments + 13 gradients + bookkeeping ~ 230 B per visible surfel).
//
// Multi-GPU (views sharded over ranks) splits the pass around the all-reduce:
//   SB_UPD_DP_PRE   activation backward + per-view weighting, written with scores and visibility into the
//                   fused [16,P] buffer that NCCL reduces;
//   SB_UPD_DP_POST  bookkeeping + Adam from the reduced buffer.
#include "prebwd_math.cuh"
#include "dp_sync.cuh"
#include <string.h>
#include "../../include/surfel_b200.h"

__global__ void sb_activate_kernel(int P, const float* __restrict__ opacity_raw, const float2* __restrict__ scaling_raw,
                                   const float4* __restrict__ rotation_raw, float* __restrict__ opacities,
                                   float2* __restrict__ scales, float4* __restrict__ rotations) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    opacities[i] = 1.0f / (1.0f + expf(-opacity_raw[i]));
    const float2 s = scaling_raw[i];
    scales[i] = make_float2(expf(s.x), expf(s.y));
    const float4 q = rotation_raw[i];
    const float den = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);  // F.normalize eps
    rotations[i] = make_float4(q.x / den, q.y / den, q.z / den, q.w / den);
}

int sb_launch_activate(int P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                       float* opacities, float* scales, float* rotations, cudaStream_t s) {
    if (P == 0) return 0;
    sb_activate_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, opacity_raw, (const float2*)scaling_raw,
                                                       (const float4*)rotation_raw, opacities, (float2*)scales,
                                                       (float4*)rotations);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

struct SbUpdateArgs {
    // raw parameters and Adam moments, group order: xyz[3] rgb[3] opacity[1] scaling[2] rotation[4]
    float* param[5];
    float* m[5];
    float* v[5];
    // gradients w.r.t. the rasterizer inputs (activated), as sb200_backward writes them
    const float* d_means3D; const float* d_colors; const float* d_opacity; const float* d_scales; const float* d_rot;
    const float* cur_scores;   // [P,2] this view's scores
    const int* radii;          // [P]
    const uint8_t* stable;     // [P] bool
    float* local_scores;       // [P,2]
    float* global_scores;      // [P,2]
    float* kf_max;             // [P]
    long long* kf_id;          // [P]
    float* fused;              // [16,P] column blocks (DP modes) or NULL
    float lr[5];
    float b1, b2, eps;
    long long kf;
    int P, mode;
};


// column c of the 13 raw-parameter columns -> (group, columns of the group, column inside the group)
#define SB_GRP(c) ((c) < 3 ? 0 : (c) < 6 ? 1 : (c) < 7 ? 2 : (c) < 9 ? 3 : 4)
#define SB_GRP_NC(g) ((g) == 0 ? 3 : (g) == 1 ? 3 : (g) == 2 ? 1 : (g) == 3 ? 2 : 4)
#define SB_GRP_C0(g) ((g) == 0 ? 0 : (g) == 1 ? 3 : (g) == 2 ? 6 : (g) == 3 ? 7 : 9)

__global__ void __launch_bounds__(256) sb_surfel_update_kernel(SbUpdateArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int P = a.P;
    if (i >= P) return;
    const size_t Pz = (size_t)P, iz = (size_t)i;
    const int mode = a.mode;

    // ---- round trip 1 (5 bytes): is this surfel updated at all? ----
    bool visible = mode == SB200_UPDATE_DP_POST ? a.fused[15 * Pz + iz] > 0.0f : a.radii[i] > 0;
    const bool adam = mode != SB200_UPDATE_DP_PRE && visible && !a.stable[i];

    // ---- round trip 2: every other load of the pass is issued before the first store (the pointers may alias
    // as far as the compiler knows, so loads never move above a store on their own; a load-update-store loop per
    // scalar was 0.20 ms at 1.1 TB/s, loads-then-stores per phase 0.17 ms) ----
    float g[13], pv[13], mv[13], vv[13];
    float cur0, cur1;
    if (mode == SB200_UPDATE_DP_POST) {
#pragma unroll
        for (int c = 0; c < 13; c++) {
            const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
            g[c] = a.fused[(size_t)SB_GRP_C0(grp) * Pz + iz * nc + (c - SB_GRP_C0(grp))];
        }
        cur0 = a.fused[13 * Pz + 2 * iz];
        cur1 = a.fused[13 * Pz + 2 * iz + 1];
    } else {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            g[k] = a.d_means3D[3 * iz + k];
            g[3 + k] = a.d_colors[3 * iz + k];
        }
        g[6] = a.d_opacity[i];
        const float2 ds = ((const float2*)a.d_scales)[i];
        g[7] = ds.x; g[8] = ds.y;
        const float4 dq = ((const float4*)a.d_rot)[i];
        g[9] = dq.x; g[10] = dq.y; g[11] = dq.z; g[12] = dq.w;
        const float2 cs = ((const float2*)a.cur_scores)[i];
        cur0 = cs.x; cur1 = cs.y;
    }
    float glob0 = a.global_scores[2 * iz], glob1 = 0.f, loc0 = 0.f, loc1 = 0.f, kfm = 0.f;
    if (mode != SB200_UPDATE_DP_PRE) {
        glob1 = a.global_scores[2 * iz + 1];
        loc0 = a.local_scores[2 * iz];
        loc1 = a.local_scores[2 * iz + 1];
        kfm = a.kf_max[i];
    }
    // raw opacity / scaling / rotation feed the activation backward of every surfel; the rest is Adam state
    const bool need_raw = mode != SB200_UPDATE_DP_POST;
#pragma unroll
    for (int c = 0; c < 7; c++) {  // xyz, rgb ([P,3] rows: scalar accesses), opacity
        const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
        const size_t e = iz * nc + (c - SB_GRP_C0(grp));
        pv[c] = (adam || (need_raw && c >= 6)) ? a.param[grp][e] : 0.f;
        mv[c] = adam ? a.m[grp][e] : 0.f;
        vv[c] = adam ? a.v[grp][e] : 0.f;
    }
    {   // scaling [P,2] and rotation [P,4] rows as one 64- / 128-bit access each (tensor bases are 16-B aligned)
        float2 s2 = make_float2(0.f, 0.f), ms2 = s2, vs2 = s2;
        float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f), mq4 = q4, vq4 = q4;
        if (adam || need_raw) {
            s2 = ((const float2*)a.param[3])[i];
            q4 = ((const float4*)a.param[4])[i];
        }
        if (adam) {
            ms2 = ((const float2*)a.m[3])[i]; vs2 = ((const float2*)a.v[3])[i];
            mq4 = ((const float4*)a.m[4])[i]; vq4 = ((const float4*)a.v[4])[i];
        }
        pv[7] = s2.x; pv[8] = s2.y; mv[7] = ms2.x; mv[8] = ms2.y; vv[7] = vs2.x; vv[8] = vs2.y;
        pv[9] = q4.x; pv[10] = q4.y; pv[11] = q4.z; pv[12] = q4.w;
        mv[9] = mq4.x; mv[10] = mq4.y; mv[11] = mq4.z; mv[12] = mq4.w;
        vv[9] = vq4.x; vv[10] = vq4.y; vv[11] = vq4.z; vv[12] = vq4.w;
    }

    if (mode != SB200_UPDATE_DP_POST) {
        // ---- backward of the activations (sigmoid, exp, normalize) ----
        const float y = 1.0f / (1.0f + expf(-pv[6]));
        g[6] = g[6] * (1.0f - y) * y;
        g[7] = g[7] * expf(pv[7]);
        g[8] = g[8] * expf(pv[8]);
        const float4 q = make_float4(pv[9], pv[10], pv[11], pv[12]);
        const float4 dq = make_float4(g[9], g[10], g[11], g[12]);
        const float len = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        if (len > 1e-12f) {
            const float4 r = make_float4(q.x / len, q.y / len, q.z / len, q.w / len);
            const float rd = r.x * dq.x + r.y * dq.y + r.z * dq.z + r.w * dq.w;
            g[9] = (dq.x - r.x * rd) / len; g[10] = (dq.y - r.y * rd) / len;
            g[11] = (dq.z - r.z * rd) / len; g[12] = (dq.w - r.w * rd) / len;
        } else {
            g[9] = dq.x / 1e-12f; g[10] = dq.y / 1e-12f; g[11] = dq.z / 1e-12f; g[12] = dq.w / 1e-12f;
        }
    }

    // ---- add_records (gaussian_model.py:163-168, gaussian_base.py:388-392) ----
    glob0 = fminf(fmaxf(glob0 + cur0, 0.0f), 1e4f);  // DP_PRE: this view's score counted, nothing stored yet
    if (mode != SB200_UPDATE_DP_PRE) {
        a.local_scores[2 * iz] = loc0 + cur0;
        if (cur1 > loc1) a.local_scores[2 * iz + 1] = cur1;
        a.global_scores[2 * iz] = glob0;
        a.global_scores[2 * iz + 1] = fminf(fmaxf(glob1, 0.0f), 1e4f);
        if (kfm < cur0) {
            a.kf_max[i] = cur0;
            a.kf_id[i] = a.kf;
        }
    }

    // ---- weighting_grad (gaussian_utils.py:56-62) ----
    if (mode != SB200_UPDATE_DP_POST) {
        const float gw = cur0 / (glob0 + 1e-6f + cur0);
#pragma unroll
        for (int c = 0; c < 13; c++) g[c] *= gw;
    }

    if (mode == SB200_UPDATE_DP_PRE) {
#pragma unroll
        for (int c = 0; c < 13; c++) {
            const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
            a.fused[(size_t)SB_GRP_C0(grp) * Pz + iz * nc + (c - SB_GRP_C0(grp))] = g[c];
        }
        a.fused[13 * Pz + 2 * iz] = cur0;
        a.fused[13 * Pz + 2 * iz + 1] = cur1;
        a.fused[15 * Pz + iz] = visible ? 1.0f : 0.0f;
        return;
    }

    // ---- masked Adam without bias correction (adam.cu:9-38) ----
    if (adam) {
        float pn[13], mn[13], vn[13];
#pragma unroll
        for (int c = 0; c < 13; c++) {
            pn[c] = pv[c]; mn[c] = mv[c]; vn[c] = vv[c];
            sb_adam_scalar(pn[c], mn[c], vn[c], g[c], a.lr[SB_GRP(c)], a.b1, a.b2, a.eps);
        }
#pragma unroll
        for (int c = 0; c < 7; c++) {
            const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
            const size_t e = iz * nc + (c - SB_GRP_C0(grp));
            a.param[grp][e] = pn[c];
            a.m[grp][e] = mn[c];
            a.v[grp][e] = vn[c];
        }
        ((float2*)a.param[3])[i] = make_float2(pn[7], pn[8]);
        ((float2*)a.m[3])[i] = make_float2(mn[7], mn[8]);
        ((float2*)a.v[3])[i] = make_float2(vn[7], vn[8]);
        ((float4*)a.param[4])[i] = make_float4(pn[9], pn[10], pn[11], pn[12]);
        ((float4*)a.m[4])[i] = make_float4(mn[9], mn[10], mn[11], mn[12]);
        ((float4*)a.v[4])[i] = make_float4(vn[9], vn[10], vn[11], vn[12]);
    }
}

int sb_launch_surfel_update(const SbUpdateArgs& a, cudaStream_t s) {
    if (a.P == 0) return 0;
    sb_surfel_update_kernel<<<(a.P + 255) / 256, 256, 0, s>>>(a);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// ---- the backward preprocess and the per-surfel update in ONE pass over the visible surfels (round 2) ----
// sb_preprocess_backward_kernel writes 72 B of dense gradient rows for EVERY surfel (zeros for the ~60 % a view does
// not see) only for sb_surfel_update_kernel to read them back: 612 MB of measured DRAM traffic per C2 view against
// ~240 MB that is compulsory (profiles/traffic.json, round 1: 1.43x and 1.59x the algorithmic bytes).  Here a thread
// looks at its surfel's radius first and leaves at once if the view did not see it (add_records with zero scores and a
// masked-out Adam step are the identity); a visible surfel's accumulated record goes through the backward preprocess
// arithmetic (prebwd_math.cuh), the activations' backward, the bookkeeping and Adam without touching memory in
// between.  MODE 0: the single-GPU update.  MODE 1: the multi-GPU stage row + visibility byte (dp_exchange.cu).
struct SbViewUpdateArgs {
    SbUpdateArgs u;                // params, moments, bookkeeping, lr, betas (gradient pointers unused)
    const float* means3D; const float* act_scales; const float* act_rot;   // the rasterizer's inputs of this view
    const float* viewmatrix; const float* projmatrix;
    float* grad_acc;               // [P, SB_GRAD_F] the backward blend's accumulation records
    int rezero;                    // hand every consumed record back zeroed (the next backward then needs no memset)
    float* stage; uint8_t* vis;    // MODE 1
    SbDpSync sync;                 // MODE 1: hand-off to the peers' reduce kernels
    int Wq, Hq;
};

template <int MODE>
__device__ __forceinline__ void sb_view_update_body(const SbViewUpdateArgs& va, const float* s_vm, const float* s_pm) {
    const SbUpdateArgs& a = va.u;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.P) return;
    const bool visible = a.radii[i] > 0;
    if (MODE == 1) va.vis[i] = visible ? 1 : 0;
    if (!visible) return;
    const size_t iz = (size_t)i;

    float acc[SB_GRAD_F];
    {
        float4* ga = reinterpret_cast<float4*>(va.grad_acc + iz * SB_GRAD_F);
#pragma unroll
        for (int k = 0; k < SB_GRAD_F / 4; k++) {
            const float4 v = ga[k];
            acc[4 * k] = v.x; acc[4 * k + 1] = v.y; acc[4 * k + 2] = v.z; acc[4 * k + 3] = v.w;
        }

    }
    float3 d_mean3D = {0.f, 0.f, 0.f};
    float2 d_scale = {0.f, 0.f};
    float4 d_rot = {0.f, 0.f, 0.f, 0.f};
    sb_prebwd_surfel<false>(i, va.Wq, va.Hq, va.means3D, (const float2*)va.act_scales, (const float4*)va.act_rot, nullptr,
                            s_vm, s_pm, acc, d_mean3D, d_scale, d_rot);
    float g[13] = {d_mean3D.x, d_mean3D.y, d_mean3D.z, acc[15], acc[16], acc[17], acc[14], d_scale.x, d_scale.y,
                   d_rot.x, d_rot.y, d_rot.z, d_rot.w};
    const float cur0 = acc[18], cur1 = acc[19];

    // raw opacity / scaling / rotation for the activations' backward; parameters and moments for Adam
    const bool adam = MODE == 0 && !a.stable[i];
    float pv[13], mv[13], vv[13];
#pragma unroll
    for (int c = 0; c < 7; c++) {
        const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
        const size_t e = iz * nc + (c - SB_GRP_C0(grp));
        pv[c] = (adam || c >= 6) ? a.param[grp][e] : 0.f;
        mv[c] = adam ? a.m[grp][e] : 0.f;
        vv[c] = adam ? a.v[grp][e] : 0.f;
    }
    {
        const float2 s2 = ((const float2*)a.param[3])[i];
        const float4 q4 = ((const float4*)a.param[4])[i];
        float2 ms2 = make_float2(0.f, 0.f), vs2 = ms2;
        float4 mq4 = make_float4(0.f, 0.f, 0.f, 0.f), vq4 = mq4;
        if (adam) {
            ms2 = ((const float2*)a.m[3])[i]; vs2 = ((const float2*)a.v[3])[i];
            mq4 = ((const float4*)a.m[4])[i]; vq4 = ((const float4*)a.v[4])[i];
        }
        pv[7] = s2.x; pv[8] = s2.y; mv[7] = ms2.x; mv[8] = ms2.y; vv[7] = vs2.x; vv[8] = vs2.y;
        pv[9] = q4.x; pv[10] = q4.y; pv[11] = q4.z; pv[12] = q4.w;
        mv[9] = mq4.x; mv[10] = mq4.y; mv[11] = mq4.z; mv[12] = mq4.w;
        vv[9] = vq4.x; vv[10] = vq4.y; vv[11] = vq4.z; vv[12] = vq4.w;
    }
    float glob0 = 0.f, glob1 = 0.f, loc0 = 0.f, loc1 = 0.f, kfm = 0.f;
    if (MODE == 0) {
        glob0 = a.global_scores[2 * iz]; glob1 = a.global_scores[2 * iz + 1];
        loc0 = a.local_scores[2 * iz]; loc1 = a.local_scores[2 * iz + 1];
        kfm = a.kf_max[i];
    }

    // Only visible surfels ever receive REDs, so zeroing what was consumed keeps the whole buffer zero: 80 B x V written
    // here instead of an 80 B x P memset in front of every backward blend.  Placed behind the last load of the pass:
    // a store in front of the parameter / moment loads serialises them (the compiler must assume aliasing) and doubled
    // this kernel's time.
    if (va.rezero) {
        float4* ga = reinterpret_cast<float4*>(va.grad_acc + iz * SB_GRAD_F);
#pragma unroll
        for (int k = 0; k < SB_GRAD_F / 4; k++) ga[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // ---- backward of the activations: the expressions of sb_surfel_update_kernel / sb_dp_stage_kernel ----
    {
        const float y = 1.0f / (1.0f + expf(-pv[6]));
        g[6] = g[6] * (1.0f - y) * y;
        g[7] = g[7] * expf(pv[7]);
        g[8] = g[8] * expf(pv[8]);
        const float4 q = make_float4(pv[9], pv[10], pv[11], pv[12]);
        const float4 dq = make_float4(g[9], g[10], g[11], g[12]);
        const float len = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        if (len > 1e-12f) {
            const float4 r = make_float4(q.x / len, q.y / len, q.z / len, q.w / len);
            const float rd = r.x * dq.x + r.y * dq.y + r.z * dq.z + r.w * dq.w;
            g[9] = (dq.x - r.x * rd) / len; g[10] = (dq.y - r.y * rd) / len;
            g[11] = (dq.z - r.z * rd) / len; g[12] = (dq.w - r.w * rd) / len;
        } else {
            g[9] = dq.x / 1e-12f; g[10] = dq.y / 1e-12f; g[11] = dq.z / 1e-12f; g[12] = dq.w / 1e-12f;
        }
    }
    if (MODE == 1) {
        float4* row = (float4*)(va.stage + iz * 16);
        row[0] = make_float4(g[0], g[1], g[2], g[3]);
        row[1] = make_float4(g[4], g[5], g[6], g[7]);
        row[2] = make_float4(g[8], g[9], g[10], g[11]);
        row[3] = make_float4(g[12], cur0, cur1, 0.0f);
        return;
    }

    // ---- add_records (gaussian_model.py:163-168, gaussian_base.py:388-392) ----
    glob0 = fminf(fmaxf(glob0 + cur0, 0.0f), 1e4f);
    a.local_scores[2 * iz] = loc0 + cur0;
    if (cur1 > loc1) a.local_scores[2 * iz + 1] = cur1;
    a.global_scores[2 * iz] = glob0;
    a.global_scores[2 * iz + 1] = fminf(fmaxf(glob1, 0.0f), 1e4f);
    if (kfm < cur0) {
        a.kf_max[i] = cur0;
        a.kf_id[i] = a.kf;
    }
    // ---- weighting_grad (gaussian_utils.py:56-62) + masked Adam (adam.cu:9-38) ----
    if (adam) {
        const float gw = cur0 / (glob0 + 1e-6f + cur0);
        float pn[13], mn[13], vn[13];
#pragma unroll
        for (int c = 0; c < 13; c++) {
            pn[c] = pv[c]; mn[c] = mv[c]; vn[c] = vv[c];
            sb_adam_scalar(pn[c], mn[c], vn[c], g[c] * gw, a.lr[SB_GRP(c)], a.b1, a.b2, a.eps);
        }
#pragma unroll
        for (int c = 0; c < 7; c++) {
            const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
            const size_t e = iz * nc + (c - SB_GRP_C0(grp));
            a.param[grp][e] = pn[c]; a.m[grp][e] = mn[c]; a.v[grp][e] = vn[c];
        }
        ((float2*)a.param[3])[i] = make_float2(pn[7], pn[8]);
        ((float2*)a.m[3])[i] = make_float2(mn[7], mn[8]);
        ((float2*)a.v[3])[i] = make_float2(vn[7], vn[8]);
        ((float4*)a.param[4])[i] = make_float4(pn[9], pn[10], pn[11], pn[12]);
        ((float4*)a.m[4])[i] = make_float4(mn[9], mn[10], mn[11], mn[12]);
        ((float4*)a.v[4])[i] = make_float4(vn[9], vn[10], vn[11], vn[12]);
    }
}

template <int MODE>
__global__ void __launch_bounds__(256) sb_view_update_kernel(SbViewUpdateArgs va) {
    __shared__ float s_vm[16], s_pm[16];
    if (threadIdx.x < 16) s_vm[threadIdx.x] = va.viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_pm[threadIdx.x - 16] = va.projmatrix[threadIdx.x - 16];
    __syncthreads();
    sb_view_update_body<MODE>(va, s_vm, s_pm);
    if (MODE == 1) sb_dp_signal(va.sync, 0);
}

int sb_dp_begin_step(void* handle, int P, float** stage, uint8_t** vis, SbDpSync* sync);   // dp_exchange.cu

int sb_launch_view_update(int mode, int W, int H, float tanfovx, float tanfovy, SbViewUpdateArgs& va, cudaStream_t s) {
    if (va.u.P == 0) return 0;
    // Q8: the reference re-derives the image size from focal*tan in fp32 (backward.cu:1022-1023)
    const float focal_y = H / (2.0f * tanfovy);
    const float focal_x = W / (2.0f * tanfovx);
    volatile float fx2 = focal_x * tanfovx;
    volatile float fy2 = focal_y * tanfovy;
    va.Wq = int(fx2 * 2);
    va.Hq = int(fy2 * 2);
    if (mode == 0) sb_view_update_kernel<0><<<(va.u.P + 255) / 256, 256, 0, s>>>(va);
    else sb_view_update_kernel<1><<<(va.u.P + 255) / 256, 256, 0, s>>>(va);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int sb200_backward_update(int P, int W, int H, int64_t R, const float* bg, float tanfovx, float tanfovy,
                                     const float* means3D, const float* scales, const float* rotations,
                                     const float* viewmatrix, const float* projmatrix, const int* radii,
                                     const void* geom_ws, const uint32_t* point_list, const void* image_ws,
                                     const float* out_color, const float* out_others, const float* dL_dout_color,
                                     const float* dL_dout_others, const uint8_t* pixel_mask, void* grad_ws, int mode,
                                     float* const* params, float* const* exp_avg, float* const* exp_avg_sq,
                                     const uint8_t* stable_mask, float* local_scores, float* global_scores,
                                     float* kf_max_scores, int64_t* kf_id, int64_t kf, const float* lr, float b1,
                                     float b2, float eps, void* dp, void* stream) {
    const bool known_zero = (mode & SB200_BWD_RECORDS_ZERO) != 0;   // grad_ws is zero on entry (left so by the previous call)
    mode &= ~SB200_BWD_RECORDS_ZERO;
    if (P < 0 || W <= 0 || H <= 0 || (mode != 0 && mode != 1)) { sb_set_error("backward_update: bad P / size / mode"); return 1; }
    if (P == 0 && mode == 1) { sb_set_error("backward_update: the multi-GPU exchange needs at least one surfel"); return 1; }
    if (P == 0) return 0;
    if (!means3D || !scales || !rotations || !viewmatrix || !projmatrix || !radii || !geom_ws || !image_ws || !grad_ws ||
        !dL_dout_color || !dL_dout_others || !pixel_mask || !params) { sb_set_error("backward_update: NULL input"); return 1; }
    if (mode == 0 && (!exp_avg || !exp_avg_sq || !stable_mask || !local_scores || !global_scores || !kf_max_scores ||
                      !kf_id || !lr)) { sb_set_error("backward_update: NULL update input"); return 1; }
    if (mode == 1 && !dp) { sb_set_error("backward_update: mode 1 needs the exchange handle"); return 1; }
    if ((((uintptr_t)params[3] | (uintptr_t)scales) & 7) || (((uintptr_t)params[4] | (uintptr_t)rotations) & 15)) {
        sb_set_error("backward_update: [P,2] arrays must be 8-byte, [P,4] arrays 16-byte aligned");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (R >= 0) {   // R < 0 (test hook): grad_ws already holds this view's accumulation records
        SbStageScope sc(SB200_STAGE_BLEND_BWD, s);
        if (!known_zero) SB_CUDA_CHECK(cudaMemsetAsync(grad_ws, 0, (size_t)P * SB_GRAD_F * 4, s));
        if (R > 0) {
            rc = sb_launch_blend_backward(W, H, bg, (const char*)geom_ws, point_list, (const char*)image_ws, out_color,
                                          out_others, dL_dout_color, dL_dout_others, (const bool*)pixel_mask,
                                          (float*)grad_ws,
                                          (const uint32_t*)((const char*)point_list + sb_pair_list_layout(R, W, H).masks),
                                          nullptr, s);
            if (rc) return rc;
        }
    }
    SbViewUpdateArgs va;
    memset(&va, 0, sizeof(va));
    for (int g = 0; g < 5; g++) {
        va.u.param[g] = params[g];
        va.u.m[g] = exp_avg ? exp_avg[g] : nullptr;
        va.u.v[g] = exp_avg_sq ? exp_avg_sq[g] : nullptr;
        va.u.lr[g] = lr ? lr[g] : 0.f;
    }
    va.u.radii = radii; va.u.stable = stable_mask; va.u.local_scores = local_scores; va.u.global_scores = global_scores;
    va.u.kf_max = kf_max_scores; va.u.kf_id = (long long*)kf_id; va.u.b1 = b1; va.u.b2 = b2; va.u.eps = eps; va.u.kf = kf;
    va.u.P = P; va.u.mode = mode;
    va.means3D = means3D; va.act_scales = scales; va.act_rot = rotations; va.viewmatrix = viewmatrix; va.projmatrix = projmatrix;
    va.grad_acc = (float*)grad_ws;
    va.rezero = R >= 0;   // (the R < 0 test hook reads records somebody else still needs)
    if (mode == 1 && sb_dp_begin_step(dp, P, &va.stage, &va.vis, &va.sync)) return 1;
    SbStageScope sc(SB200_STAGE_UPDATE, s);
    return sb_launch_view_update(mode, W, H, tanfovx, tanfovy, va, s);
}

extern "C" int sb200_surfel_update(int P, int mode, float* const* params, float* const* exp_avg,
                                   float* const* exp_avg_sq, const float* d_means3D, const float* d_colors,
                                   const float* d_opacity, const float* d_scales, const float* d_rotations,
                                   const float* cur_scores, const int* radii, const uint8_t* stable_mask,
                                   float* local_scores, float* global_scores, float* kf_max_scores, int64_t* kf_id,
                                   int64_t kf, float* fused, const float* lr, float b1, float b2, float eps,
                                   void* stream) {
    if (P < 0 || mode < 0 || mode > 2) { sb_set_error("surfel_update: bad P/mode"); return 1; }
    if (!params || !exp_avg || !exp_avg_sq || !lr || !global_scores) { sb_set_error("surfel_update: NULL input"); return 1; }
    if (mode != SB200_UPDATE_LOCAL && !fused) { sb_set_error("surfel_update: DP modes need the fused buffer"); return 1; }
    if (mode != SB200_UPDATE_DP_POST &&
        (!d_means3D || !d_colors || !d_opacity || !d_scales || !d_rotations || !cur_scores || !radii)) {
        sb_set_error("surfel_update: NULL gradient input");
        return 1;
    }
    if (mode != SB200_UPDATE_DP_PRE && (!stable_mask || !local_scores || !kf_max_scores || !kf_id)) {
        sb_set_error("surfel_update: NULL bookkeeping input");
        return 1;
    }
    // [P,2] and [P,4] rows are accessed as float2 / float4
    if ((((uintptr_t)params[3] | (uintptr_t)exp_avg[3] | (uintptr_t)exp_avg_sq[3] | (uintptr_t)d_scales |
          (uintptr_t)cur_scores) & 7) ||
        (((uintptr_t)params[4] | (uintptr_t)exp_avg[4] | (uintptr_t)exp_avg_sq[4] | (uintptr_t)d_rotations) & 15)) {
        sb_set_error("surfel_update: scaling / score arrays must be 8-byte, rotation arrays 16-byte aligned");
        return 1;
    }
    SbUpdateArgs a;
    for (int g = 0; g < 5; g++) {
        a.param[g] = params[g]; a.m[g] = exp_avg[g]; a.v[g] = exp_avg_sq[g]; a.lr[g] = lr[g];
    }
    a.d_means3D = d_means3D; a.d_colors = d_colors; a.d_opacity = d_opacity; a.d_scales = d_scales; a.d_rot = d_rotations;
    a.cur_scores = cur_scores; a.radii = radii; a.stable = stable_mask;
    a.local_scores = local_scores; a.global_scores = global_scores; a.kf_max = kf_max_scores; a.kf_id = (long long*)kf_id;
    a.fused = fused; a.b1 = b1; a.b2 = b2; a.eps = eps; a.kf = kf; a.P = P; a.mode = mode;
    SbStageScope sc(SB200_STAGE_UPDATE, (cudaStream_t)stream);
    return sb_launch_surfel_update(a, (cudaStream_t)stream);
}
