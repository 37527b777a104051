// K4a: backward of the per-tile blend.
//
// Restates the *live* backward of the reference, PerGaussianRenderCUDA
// (DSR/cuda_rasterizer/backward.cu:445-845), including its quirks:
// the gradient formulas use prefix-minus-final accumulators "ar", a transmittance re-propagated front
// to back, M1/M2 frozen at each 32-splat bucket start, and no p.z==0 / depth<near skips.
//
// The reference needs 12-float-per-pixel checkpoints every 32 splats (written by its forward) so that
// one warp per bucket can start mid-tile.  Here one CTA owns a tile and walks it front to back once,
// recomputing the forward state on the fly, so no checkpoint is ever written or read.
//
// Structure (round 2 of this kernel; profiles/r01_blend_backward_lines.txt showed 25 % of the stall samples
// on the two CTA barriers per 32-surfel batch of the first version, with 59 % issue utilisation):
//   * 8 warps (one 8x4 pixel block each, a thread owns a pixel) that never meet at a CTA barrier inside the walk;
//   * the sorted surfels' 128-byte records are gathered (ring.cuh: 16-byte cp.async, or one TMA bulk copy per record
//     with SB_RING_LDGSTS=0) into an 8-stage shared-memory ring of 32-surfel batches; a stage is handed to the warps
//     by an mbarrier and handed back through a shared counter: the LAST warp to finish a batch refills its stage with
//     the batch 8 ahead, so a warp waits only for data it needs and the warps drift up to 7 batches apart (a dedicated
//     producer warp made 9 warps per CTA, which the four 16 K-register SM sub-partitions cannot hold twice at 112
//     registers: measured 1 CTA/SM and 3.0 ms);
//   * per hit surfel a 21-shuffle multi-value butterfly reduces the 20 gradient terms over the warp's 32
//     pixels and 20 lanes issue ONE coalesced red.global.add.f32 over the surfel's 80-byte accumulation
//     record (3 sectors); there is no cross-warp combine and no shared accumulation buffer;
//   * every warp stops at ITS OWN last contributor (warp max), not the tile's;
//   * round 2: NO culling here.  The forward blend already knows, per (32-surfel batch, warp), which surfels some live
//     pixel of the warp's block may blend, and leaves that as one 32-bit hit mask per (batch, warp) behind the sorted
//     id list (sb_pair_list_layout, common.cuh).  The ring's fill stages the batch's 8 mask words next to the ids,
//     and a warp visits exactly the set bits: the box test, the two-at-a-time conic cull and their 3 LDS.128 per
//     candidate (18 % of this kernel's instructions in round 1, profiles/r01_final_blend_backward_lines.txt) are
//     gone.  A pixel joins a visit iff the surfel lies before its last contributor; every gate is then the
//     reference's own arithmetic, so a pixel outside the surfel's footprint fails alpha >= 1/255 exactly as it does
//     in the reference, which evaluates all 256 pixels of a tile for every surfel.  The ring's fill in turn gathers
//     only the records of surfels whose bit is set in SOME warp's word (BB_SPARSE_FILL).
#include "pair_math.cuh"
#include "ring.cuh"
#include <mutex>

#define BB_CWARPS 8
#define BB_THREADS (BB_CWARPS * 32)
#define BB_BATCH 32  // == the reference's bucket size; bucket boundaries are part of the semantics (Q5)
#ifndef BB_STAGES
#define BB_STAGES 8
#endif

// Gradient-side arithmetic is not gated (only k, l, p, s, rho, G, alpha decide which pairs blend, and those stay
// bit-pinned in pair_math.cuh), so its divisions use the MUFU reciprocal: the gradients are compared with the
// reference at 5e-4 relative L1 (atomic-order jitter), fp32 approximate division is ~1e-7.
__device__ __forceinline__ float bb_ndc_depth_fast(float depth) {
    return (1.0f - SB_NEAR * __fdividef(1.0f, depth)) * (SB_FAR / (SB_FAR - SB_NEAR));
}

// Multi-value warp reduction: every lane holds v[0..K); afterwards lane L holds in v[0] the warp-wide
// sum of value slot(L) (or junk if slot(L) < 0).  Each step halves the number of live values.
template <int N, int OFF>
struct Butterfly {
    static __device__ __forceinline__ void run(float* v, int lane) {
        constexpr int HALF = (N + 1) / 2;
        const bool upper = (lane & OFF) != 0;
#pragma unroll
        for (int i = 0; i < HALF; i++) {
            float lo = v[i];
            float hi = (i + HALF < N) ? v[i + HALF] : 0.f;
            float send = upper ? lo : hi;
            float keep = upper ? hi : lo;
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
        }
        Butterfly<HALF, OFF / 2>::run(v, lane);
    }
};
template <int N>
struct Butterfly<N, 0> {
    static __device__ __forceinline__ void run(float*, int) {}
};

// Which value slot ends up in this lane's v[0] (-1: none).  n_ct follows the compile-time halving
// path of Butterfly<K,16>; n_real counts how many real (non-padding) values the lane's group holds.
__device__ __forceinline__ int butterfly_slot(int lane, int K) {
    int base = 0, n_ct = K, n_real = K;
    for (int off = 16; off > 0; off >>= 1) {
        int half = (n_ct + 1) / 2;
        if (lane & off) {
            base += half;
            n_real -= half;
        } else {
            n_real = min(n_real, half);
        }
        n_ct = half;
    }
    return n_real >= 1 ? base : -1;
}

// Occupancy against spills (measured): 3 CTAs/SM = 80 registers, 4 CTAs/SM = 64 registers.  On the
// KITTI-shaped image (1872 tiles) 3 is 5 % faster, on the 1080p-class images (8160 / 9600 tiles, deep lists) 4 is 7-9 %
// faster; 2 CTAs/SM (112 registers, no spills) loses everywhere.  Both are instantiated and the launcher picks by tile
// count.
//
// What ptxas spilled were the per-pixel CONSTANTS of the gradient arithmetic (the pixel's upstream gradients and final
// T / M1 / distortion): 6 local-memory loads per blended pair and one per bucket start, the latter alone 13 % of the
// kernel's stall samples (profiles/r02_blend_backward_lines.txt, line `ar_reg = reg_f - fin_reg`).  They now live in
// shared memory, [k][thread] (conflict-free), read once per use: all 13 of them at 3 CTAs/SM (spills 72 -> 12 B, none
// left in the per-pair path; -3.7 % at C2), the first 5 at 4 CTAs/SM, where shared memory is what limits the L1 that
// catches the remaining spills (5: -3.3 % at C4 / C5; 7 or 10: +11..17 %).  profiles/r02_ab_variants.txt, blocks r02j-l.
#ifndef BB_PIX_N3
#define BB_PIX_N3 13
#endif
#ifndef BB_PIX_N4
#define BB_PIX_N4 5
#endif
// The ring gathers only the records some warp of the tile will visit (the union of the batch's 8 hit masks).  Measured
// (profiles/r02_ab_variants.txt, block r02m): C2 1.143 -> 1.133 ms, C4 3.90 -> 3.17, C5 5.28 -> 4.03 -- on the
// 1080p-class configs 4 of 5 list entries are surfels whose footprint misses the whole tile.
#ifndef BB_SPARSE_FILL
#define BB_SPARSE_FILL 1
#endif
#ifndef BB_DEEP_TILES
#define BB_DEEP_TILES 4000
#endif
template <int MIN_CTAS>
__global__ void __launch_bounds__(BB_THREADS, MIN_CTAS)
sb_blend_backward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                         const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
                         const float* __restrict__ bg, const float* __restrict__ final_T,
                         const uint32_t* __restrict__ n_contrib, const float* __restrict__ out_color,
                         const float* __restrict__ out_others, const float* __restrict__ dL_dpixels,
                         const float* __restrict__ dL_dothers, const bool* __restrict__ pixel_mask,
                         float* __restrict__ grad_acc, const uint32_t* __restrict__ hit_masks,
                         const uint32_t* __restrict__ walk_cap) {
    __shared__ __align__(128) float4 s_rec[BB_STAGES][BB_BATCH][SB_REC_F4];
    __shared__ uint32_t s_id[BB_STAGES][BB_BATCH];
    __shared__ uint32_t s_mask[BB_STAGES][BB_CWARPS];
    __shared__ __align__(8) uint64_t s_full[BB_STAGES];
    __shared__ uint32_t s_done[BB_STAGES];
    __shared__ uint32_t s_maxlast[BB_CWARPS];
    __shared__ float s_shadow[9][BB_THREADS];  // forward-state shadow of the Q4 case: ar[8], T
    // per-pixel constants of the gradient arithmetic (see above): 0..2 dLn, 3 M1_final, 4 fin_reg, 5 bg . dL_dpixel,
    // 6 dLa, 7 dLd, 8 dLreg, 9 T_final, 10..12 dLc; the first PIX_N of them live here, the rest stay in registers
    constexpr int PIX_N = MIN_CTAS >= 4 ? BB_PIX_N4 : BB_PIX_N3;
    extern __shared__ float s_pix_dyn[];
    float(*s_pix)[BB_THREADS] = (float(*)[BB_THREADS])s_pix_dyn;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
#ifdef SB_DEBUG_NO_LPT
    const uint32_t tile = blockIdx.x;
#else
    const uint32_t tile = tile_order[blockIdx.x];  // longest lists first (binning.cu)
#endif
    const uint32_t tx = tile % grid_x, ty = tile / grid_x;
    // warp w owns the 8x4 block at (w&1, w>>1); lane l the pixel (l&7, l>>3) of it (as in the forward)
    const uint32_t wx0 = tx * SB_TILE_X + (wid & 1) * 8, wy0 = ty * SB_TILE_Y + (wid >> 1) * 4;
    const uint32_t px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float2 pixf = {(float)px, (float)py};
    const size_t HW = (size_t)H * W;

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);

    // per-pixel constants
    uint32_t last = 0;
    float T_final = 0.f, M1_final = 0.f;
    float dLc[3] = {0.f, 0.f, 0.f}, dLn[3] = {0.f, 0.f, 0.f};
    float dLd = 0.f, dLa = 0.f, dLreg = 0.f, fin_reg = 0.f;
    // "ar" accumulators: prefix - final  (colour 0..2, depth 3, accum 4, normal 5..7)
    float ar[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (inside && pixel_mask[pix_id]) {  // masked pixels contribute nothing (Q14)
        last = n_contrib[pix_id];
        T_final = final_T[pix_id];
        M1_final = final_T[pix_id + HW];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
            dLc[ch] = dL_dpixels[ch * HW + pix_id];
            ar[ch] = -out_color[ch * HW + pix_id];  // includes T_final*bg, as in the reference (Q6)
            dLn[ch] = dL_dothers[(2 + ch) * HW + pix_id];
            ar[5 + ch] = -out_others[(2 + ch) * HW + pix_id];
        }
        dLd = dL_dothers[pix_id];
        dLa = dL_dothers[HW + pix_id];
        dLreg = dL_dothers[6 * HW + pix_id];
        ar[3] = -out_others[pix_id];
        ar[4] = -out_others[HW + pix_id];
        fin_reg = out_others[6 * HW + pix_id];
        // median-depth channel (5) deliberately receives no gradient (Q3)
    }
    // sum_ch bg[ch] * dL_dpixel[ch], accumulated in the reference's order (backward.cu:712-718); constant per pixel
    float bg_dot_dpixel = 0.0f;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) bg_dot_dpixel += bg[ch] * dLc[ch];
    {
        const float pv[13] = {dLn[0], dLn[1], dLn[2], M1_final, fin_reg, bg_dot_dpixel, dLa, dLd, dLreg, T_final,
                              dLc[0], dLc[1], dLc[2]};
#pragma unroll
        for (int k = 0; k < PIX_N; k++) s_pix[k][tid] = pv[k];
    }
    const volatile float* pixc = &s_pix[0][tid];   // volatile: one LDS per use below, nothing hoisted back into registers
#define BB_PIX(k, reg) ((k) < PIX_N ? pixc[(k) * BB_THREADS] : (reg))

    // tile-wide max of `last` bounds the walk (reference: max_contrib bucket skip, backward.cu:503-505);
    // each warp blends only up to its own max
    uint32_t wmax = last;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) s_maxlast[wid] = wmax;
    if (tid < BB_STAGES) {
        bb_mbar_init(&s_full[tid], BB_RING_ARRIVALS);
        s_done[tid] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t maxlast = 0;
#pragma unroll
    for (int w = 0; w < BB_CWARPS; w++) maxlast = max(maxlast, s_maxlast[w]);
    int n_walk = min(n, (int)maxlast);
    // test hook (sb200_debug_walk_cap): stop where the reference's stored max_contrib makes ITS backward stop
    // (whole 32-surfel buckets at or beyond it are dropped, backward.cu:503-505; quirk Q17 in DESIGN.md)
    // (the stored value can be stale shared memory, i.e. any 32-bit pattern: clamp before rounding up)
    if (walk_cap) n_walk = min(n_walk, (int)((min(walk_cap[tile], 0x40000000u) + BB_BATCH - 1) / BB_BATCH) * BB_BATCH);
    const int nbatches = (n_walk + BB_BATCH - 1) / BB_BATCH;

    const uint32_t* tile_masks = hit_masks + (size_t)sb_hit_slot(range.x, tile) * BB_CWARPS;
    // gather the records of batch b into ring stage b % BB_STAGES (called by one whole warp)
    auto fill = [&](int b) {
        const int st = b % BB_STAGES;
        const int cnt = min(BB_BATCH, n_walk - b * BB_BATCH);
        uint32_t mw = 0;
        if (lane < BB_CWARPS) {
            mw = __ldg(tile_masks + (size_t)b * BB_CWARPS + lane);
            s_mask[st][lane] = mw;
        }
#if BB_SPARSE_FILL
        // only the surfels some warp of the tile will visit are gathered (the union of the 8 hit masks): on the
        // 1080p-class configs 4 of 5 list entries are surfels whose footprint misses the whole tile
        const uint32_t want = __reduce_or_sync(0xffffffffu, mw) & (cnt < BB_BATCH ? (1u << cnt) - 1u : 0xffffffffu);
#else
        const uint32_t want = cnt < BB_BATCH ? (1u << cnt) - 1u : 0xffffffffu;
#endif
        uint32_t id = 0;
        if ((want >> lane) & 1u) {
            id = __ldg(point_list + range.x + b * BB_BATCH + lane);
            s_id[st][lane] = id;
        }
        __syncwarp();
        bb_gather<SB_REC_F4>(s_rec[st], rec, id, want, &s_full[st], lane);
    };
    for (int b0 = wid; b0 < BB_STAGES && b0 < nbatches; b0 += BB_CWARPS) fill(b0);  // prime the ring
    const int my_walk = min(n_walk, (int)wmax);   // this warp's pixels blend nothing at or behind this list position
    const int my_batches = (my_walk + BB_BATCH - 1) / BB_BATCH;

    // running state
    float T = 1.f;              // transmittance as the backward propagates it
    float ar_reg = 0.f;         // distortion accumulator (reset from the true prefix at bucket starts)
    float M1 = 0.f, M2 = 0.f;   // true forward moments (running)
    float reg_f = 0.f;          // true forward distortion prefix
    float M1b = 0.f, M2b = 0.f; // moments frozen at the bucket start (Q5)
    // shadow of the true forward state, live only after a forward-skipped splat was blended by the
    // backward inside the current bucket (Q4); the next bucket start snaps back to it
    bool div = false;
    // (T and ar[8] of the shadow live in shared memory, s_shadow[.][tid]: touched only after a Q4 pair)

    int slot = butterfly_slot(lane, SB_GRAD_F);
    asm volatile("" : "+r"(slot));  // keep it in a register: ptxas otherwise recomputes the 5-step chain per hit (4 % of all instructions)

    for (int b = 0; b < nbatches; b++) {
        const int st = b % BB_STAGES;
        bb_mbar_wait(&s_full[st], (uint32_t)(b / BB_STAGES) & 1u);
        if (b < my_batches) {
            // ---- bucket start: what the reference reloads from its checkpoint ----
            if (div) {
                T = s_shadow[8][tid];
#pragma unroll
                for (int i = 0; i < 8; i++) ar[i] = s_shadow[i][tid];
                div = false;
            }
            ar_reg = reg_f - BB_PIX(4, fin_reg);
            M1b = M1;
            M2b = M2;
            const float4(*sr)[SB_REC_F4] = s_rec[st];
            // the surfels of this batch the forward marked for this warp, cut where the warp's walk ends: behind the
            // warp's last contributor no pixel of the block is live (and a capped walk -- sb200_debug_walk_cap --
            // ends on a batch boundary).  Below that limit every visit has at least one live pixel.
            uint32_t am = s_mask[st][wid];
            const int lim = my_walk - b * BB_BATCH;   // > 0 here
            if (lim < BB_BATCH) am &= (1u << lim) - 1u;
            while (am) {
                const int j = __ffs(am) - 1;
                am &= am - 1;
                const bool pass = (uint32_t)(b * BB_BATCH + j) < last;
                const float2 d = {sr[j][0].w - pixf.x, sr[j][1].w - pixf.y};
                float g[SB_GRAD_F];
#pragma unroll
                for (int i = 0; i < SB_GRAD_F; i++) g[i] = 0.f;
                bool hit = false;
                if (pass) {
                    const float4 q0 = sr[j][SB_REC_TU], q1 = sr[j][SB_REC_TV], q2 = sr[j][SB_REC_TW];
                    const float3 Tu = {q0.x, q0.y, q0.z}, Tv = {q1.x, q1.y, q1.z}, Tw = {q2.x, q2.y, q2.z};
                    const float opa = q0.w;
                    // pinned geometry (pair_math.cuh): the backward's own rho2d rounding for its gates,
                    // the forward's for what the forward decided
                    const SbPair pg = sb_pair(pixf, Tu, Tv, Tw);
                    const float3 k = pg.k, l = pg.l, p = pg.p;
                    const float2 s = pg.s;
                    const float rho3d = pg.rho3d;
                    const float rho2d = sb_rho2d_backward(d);
                    const float c_d = sb_pair_depth(pg, rho2d, Tw.z);
                    const float c_d_fwd = sb_pair_depth(pg, sb_rho2d_forward(d), Tw.z);
                    const float power = __fmul_rn(fminf(rho3d, rho2d), -0.5f);
                    float G = 0.f, alpha = 0.f;
                    bool hit_b = false;  // the reference's backward blends this pair (backward.cu:688-694)
                    if (!(power > 0.0f)) {
                        G = expf(power);
                        alpha = fminf(0.99f, __fmul_rn(G, opa));
                        hit_b = !(alpha < 1.0f / 255.0f);
                    }
                    // did the reference's FORWARD blend it (forward.cu:408-435)?  Same gates plus the
                    // p.z == 0 / depth < near skips the backward lacks (Q4), with the forward's rho2d rounding.
                    const float rho2d_f = sb_rho2d_forward(d);
                    float alpha_f = alpha;
                    bool contrib_f = hit_b;
                    if (rho2d_f != rho2d) {  // the two contractions disagree by an ulp: re-evaluate the gate
                        const float pf = __fmul_rn(fminf(rho3d, rho2d_f), -0.5f);
                        contrib_f = false;
                        if (!(pf > 0.0f)) {
                            alpha_f = fminf(0.99f, __fmul_rn(expf(pf), opa));
                            contrib_f = !(alpha_f < 1.0f / 255.0f);
                        }
                    }
                    contrib_f = contrib_f && !(p.z == 0.0f) && !(c_d_fwd < SB_NEAR);
                    if (hit_b || contrib_f) {
                        const float4 q6 = sr[j][SB_REC_N];
                        const float c[3] = {q1.w, q2.w, q6.w};
                        const float nrm[3] = {q6.x, q6.y, q6.z};
                        // the backward's state leaves the forward's until the next bucket start
                        if (hit_b != contrib_f && !div) {
                            div = true;
                            s_shadow[8][tid] = T;
#pragma unroll
                            for (int i = 0; i < 8; i++) s_shadow[i][tid] = ar[i];
                        }
                        const float m_d = bb_ndc_depth_fast(c_d);
                        // ---- true forward bookkeeping (moments + distortion prefix) ----
                        if (contrib_f) {
                            const float m_f = (c_d_fwd == c_d) ? m_d : bb_ndc_depth_fast(c_d_fwd);
                            const float Tt = div ? s_shadow[8][tid] : T;
                            const float wf = alpha_f * Tt;
                            reg_f += (m_f * m_f * (1 - Tt) + M2 - 2 * m_f * M1) * wf;
                            M1 += m_f * wf;
                            M2 += m_f * m_f * wf;
                            if (div) {
                                s_shadow[0][tid] += wf * c[0]; s_shadow[1][tid] += wf * c[1]; s_shadow[2][tid] += wf * c[2];
                                s_shadow[3][tid] += wf * c_d_fwd; s_shadow[4][tid] += wf;
                                s_shadow[5][tid] += wf * nrm[0]; s_shadow[6][tid] += wf * nrm[1]; s_shadow[7][tid] += wf * nrm[2];
                                s_shadow[8][tid] = Tt * (1 - alpha_f);
                            }
                        }
                        if (hit_b) {
                            hit = true;
                            const float dmd_dd = __fdividef((SB_FAR * SB_NEAR) / (SB_FAR - SB_NEAR), c_d * c_d);
                            // ---- the reference's gradient arithmetic (backward.cu:696-814) ----
                            const float dLn_[3] = {BB_PIX(0, dLn[0]), BB_PIX(1, dLn[1]), BB_PIX(2, dLn[2])};
                            const float dLc_[3] = {BB_PIX(10, dLc[0]), BB_PIX(11, dLc[1]), BB_PIX(12, dLc[2])};
                            const float dLd_ = BB_PIX(7, dLd), dLa_ = BB_PIX(6, dLa), dLreg_ = BB_PIX(8, dLreg);
                            const float T_final_ = BB_PIX(9, T_final), M1_final_ = BB_PIX(3, M1_final);
                            const float bg_dot_ = BB_PIX(5, bg_dot_dpixel);
                            const float dchannel_dcolor = alpha * T;
                            const float inv1ma = __fdividef(1.0f, 1.0f - alpha);  // alpha <= 0.99
                            float dL_dalpha = 0.0f;
#pragma unroll
                            for (int ch = 0; ch < 3; ch++) {
                                ar[ch] += T * alpha * c[ch];
                                g[15 + ch] = dchannel_dcolor * dLc_[ch];
                                dL_dalpha += ((c[ch] * T) - inv1ma * (-ar[ch])) * dLc_[ch];
                                g[19] += fabsf(dchannel_dcolor * dLc_[ch]);
                            }
                            dL_dalpha += (-T_final_ * inv1ma) * bg_dot_;
                            g[18] = dchannel_dcolor;
                            float dL_dz = 0.0f;
                            ar[3] += T * alpha * c_d;
                            dL_dalpha += ((c_d * T) - inv1ma * (-ar[3])) * dLd_;
                            dL_dz += alpha * T * dLd_;
                            ar[4] += T * alpha * 1.0f;
                            dL_dalpha += ((1.0f * T) - inv1ma * (-ar[4])) * dLa_;
#pragma unroll
                            for (int ch = 0; ch < 3; ch++) {
                                ar[5 + ch] += T * alpha * nrm[ch];
                                g[11 + ch] = dchannel_dcolor * dLn_[ch];
                                dL_dalpha += ((nrm[ch] * T) - inv1ma * (-ar[5 + ch])) * dLn_[ch];
                            }
                            const float A = 1 - T;
                            const float dist = (m_d * m_d * A + M2b - 2 * m_d * M1b);
                            ar_reg += T * alpha * dist;
                            dL_dalpha += ((dist * T) - inv1ma * (-ar_reg)) * dLreg_;
                            const float final_A = 1 - T_final_;
                            const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - M1_final_) * dLreg_;
                            dL_dz += dL_dmd * dmd_dd;
                            T *= (1.0f - alpha);

                            const float dL_dG = opa * dL_dalpha;
                            if (rho3d <= rho2d) {
                                const float2 dL_ds = {dL_dG * -G * s.x + dL_dz * Tw.x, dL_dG * -G * s.y + dL_dz * Tw.y};
                                const float inv_pz = __fdividef(1.0f, p.z);
                                const float dsx_pz = dL_ds.x * inv_pz;
                                const float dsy_pz = dL_ds.y * inv_pz;
                                const float3 dL_dp = {dsx_pz, dsy_pz, -(dsx_pz * s.x + dsy_pz * s.y)};
                                const float3 dL_dk = {l.y * dL_dp.z - l.z * dL_dp.y, l.z * dL_dp.x - l.x * dL_dp.z,
                                                      l.x * dL_dp.y - l.y * dL_dp.x};
                                const float3 dL_dl = {dL_dp.y * k.z - dL_dp.z * k.y, dL_dp.z * k.x - dL_dp.x * k.z,
                                                      dL_dp.x * k.y - dL_dp.y * k.x};
                                g[0] = -dL_dk.x; g[1] = -dL_dk.y; g[2] = -dL_dk.z;
                                g[3] = -dL_dl.x; g[4] = -dL_dl.y; g[5] = -dL_dl.z;
                                g[6] = pixf.x * dL_dk.x + pixf.y * dL_dl.x + dL_dz * s.x;
                                g[7] = pixf.x * dL_dk.y + pixf.y * dL_dl.y + dL_dz * s.y;
                                g[8] = pixf.x * dL_dk.z + pixf.y * dL_dl.z + dL_dz * 1.0f;
                            } else {
                                const float dG_ddelx = -G * SB_FILTER_INV_SQUARE * d.x;
                                const float dG_ddely = -G * SB_FILTER_INV_SQUARE * d.y;
                                g[9] = dL_dG * dG_ddelx;
                                g[10] = dL_dG * dG_ddely;
                                g[8] = dL_dz;
                            }
                            g[14] = G * dL_dalpha;
                        }
                    }
                }
                if (__any_sync(0xffffffffu, hit)) {
                    Butterfly<SB_GRAD_F, 16>::run(g, lane);
                    // 20 lanes, 20 consecutive floats of the surfel's accumulation record: one coalesced RED
                    if (slot >= 0)
                        atomicAdd(grad_acc + (size_t)s_id[st][j] * SB_GRAD_F + slot, g[0]);
                }
            }
        }
        // hand the stage back: the last of the 8 warps to get here refills it with the batch BB_STAGES ahead
        __syncwarp();
        uint32_t old = 0;
        if (lane == 0) {
            __threadfence_block();
            old = atomicAdd(&s_done[st], 1u);
        }
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old == BB_CWARPS - 1) {
            if (lane == 0) s_done[st] = 0;
            __threadfence_block();
            if (b + BB_STAGES < nbatches) fill(b + BB_STAGES);
        }
    }
    bb_ring_drain();
}

int sb_launch_blend_backward(int W, int H, const float* bg, const char* geom, const uint32_t* point_list,
                             const char* image, const float* out_color, const float* out_others,
                             const float* dL_dcolor, const float* dL_dothers, const bool* pixel_mask,
                             float* grad_acc, const uint32_t* hit_masks, const uint32_t* walk_cap, cudaStream_t s) {
    SbImageLayout IL = sb_image_layout(W, H);
    const int grid_x = (W + SB_TILE_X - 1) / SB_TILE_X, grid_y = (H + SB_TILE_Y - 1) / SB_TILE_Y;
    const bool deep = grid_x * grid_y >= BB_DEEP_TILES;
    const size_t dyn = (size_t)(deep ? BB_PIX_N4 : BB_PIX_N3) * BB_THREADS * sizeof(float);
    {   // static + dynamic shared memory passes 48 KB: opt in, once per device (function attributes are per device)
        static std::mutex mu;
        static bool attr_set[SB_MAX_DEVICES] = {};
        const int dev = sb_current_device();
        std::lock_guard<std::mutex> lock(mu);
        if (!attr_set[dev]) {
            SB_CUDA_CHECK(cudaFuncSetAttribute(sb_blend_backward_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               BB_PIX_N3 * BB_THREADS * (int)sizeof(float)));
            SB_CUDA_CHECK(cudaFuncSetAttribute(sb_blend_backward_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               BB_PIX_N4 * BB_THREADS * (int)sizeof(float)));
            attr_set[dev] = true;
        }
    }
    auto kern = deep ? sb_blend_backward_kernel<4> : sb_blend_backward_kernel<3>;
    kern<<<grid_x * grid_y, BB_THREADS, dyn, s>>>(
        W, H, grid_x, (const uint2*)(image + IL.ranges), (const uint32_t*)(image + IL.order), point_list, (const float4*)geom, bg,
        (const float*)(image + IL.final_T), (const uint32_t*)(image + IL.n_contrib), out_color, out_others,
        dL_dcolor, dL_dothers, pixel_mask, grad_acc, hit_masks, walk_cap);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
