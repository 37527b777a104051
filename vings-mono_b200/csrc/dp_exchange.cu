// Multi-GPU exchange of the mapping step, written against NVLink peer memory.
//
// The keyframe window's views are partitioned over the ranks of one box, surfels and Adam state are replicated.
// Every optimizer step each rank has rendered ONE view; per surfel the step needs the sum over the ranks' views of
// the (score-weighted) raw-parameter gradients, the score bookkeeping of every view and the union of the
// visibilities.  Round 1 did that with a dense NCCL all-reduce of a [16,P] fp32 buffer (64 B per surfel per rank,
// 60 % of it zeros: a view sees ~40 % of the cloud), fully exposed between two per-surfel kernels.
//
// Here the exchange is three kernels over peer-mapped ("symmetric") buffers -- the compute and the collective are
// the same launches, only rows of surfels a view actually saw ever cross NVLink, and no reduction order depends on
// the rank that reads the result:
//
//   sb_dp_stage_kernel   (every rank, all surfels)   backward of the three activations; a visible surfel's 13
//                        raw-parameter gradients + its two scores go into the rank's OWN stage row (64 B), its
//                        visibility into a byte array.  Local stores only.
//   -- hand-off: the LAST CTA of the stage kernel release-stores the step number into every peer's flag block; every
//      CTA of the next kernel spins (acquire, system scope) on its own rank's flags before it touches peer rows --
//   sb_dp_reduce_kernel  (rank r, the 256-surfel chunks c with c % N == r)   a reduce-scatter by peer LOADS: for the
//                        surfels it owns the rank reads the stage rows of the ranks that saw them, applies weighting_grad per
//                        view (scripts/gaussian/gaussian_utils.py:56-62) and folds the views IN RANK ORDER into one
//                        64-B result row: gradient sum, add_records' running quantities (gaussian_model.py:163-168:
//                        score sum, max of the second score, max of the first score; the rank of its first arg-max
//                        rides in the touched byte).
//   -- same hand-off on a second flag channel --
//   sb_dp_post_kernel    (every rank, all surfels)   an all-gather fused with its consumer: reads the result row of
//                        every touched surfel from its owner (peer load for 7/8 of them) and applies the score
//                        bookkeeping and the masked Adam step to the local replica.  Every rank consumes the same
//                        65 bytes, so the replicas stay bit-identical by construction.
//
// There is no barrier kernel and no collective call: the only cross-GPU synchronisation is those two flag hand-offs
// (monotonic step counters, never reset), and they also order the buffers' reuse -- a rank's next stage kernel
// follows its own post kernel, which has waited for every peer's reduce (the readers of its stage rows); its next
// reduce waits for every peer's next stage, which follows that peer's post (the readers of its result rows).
//
// Bytes over NVLink per rank and step at C2 on 8 GPUs: stage rows 7/8 * (P/8) * 8 * 0.41 * 64 B = 22 MB in, result
// rows 7/8 * ~0.6 P * 64 B = 40 MB in, against 2 * 7/8 * 77 MB = 134 MB for the dense all-reduce.
#include "dp_sync.cuh"
#include <string.h>
#include "../../include/surfel_b200.h"

#define SB_DP_STAGE_F 16   // floats per stage row: 13 gradients, score 0, score 1, pad
#define SB_DP_OUT_F 16     // floats per result row: 13 gradients, score-0 sum, score-1 max, score-0 max; the rank whose view
                           // holds that maximum travels in the `upd` byte (1 + rank; 0 = no view saw the surfel)

// Loads of rows another GPU wrote: the writer's kernel released them with a system-scope store of the step number,
// this kernel acquired that flag (sb_dp_wait) before any thread gets here, so ordinary loads observe them; .cg keeps
// them out of L1 (they are read once).
__device__ __forceinline__ float4 sb_ld_peer_f4(const float4* p) { return __ldcg(p); }
__device__ __forceinline__ uint32_t sb_ld_peer_u8(const uint8_t* p) { return (uint32_t)__ldcg(p); }

struct SbDpStageArgs {
    const float* d_means3D; const float* d_colors; const float* d_opacity; const float* d_scales; const float* d_rot;
    const float* cur_scores; const int* radii;
    const float* opacity_raw; const float* scaling_raw; const float* rotation_raw;
    float* stage; uint8_t* vis;
    int P;
    SbDpSync sync;
};

__device__ __forceinline__ void sb_dp_stage_body(const SbDpStageArgs& a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.P) return;
    const bool visible = a.radii[i] > 0;
    a.vis[i] = visible ? 1 : 0;
    if (!visible) return;
    const size_t iz = (size_t)i;
    float g[13];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        g[k] = a.d_means3D[3 * iz + k];
        g[3 + k] = a.d_colors[3 * iz + k];
    }
    g[6] = a.d_opacity[i];
    const float2 ds = ((const float2*)a.d_scales)[i];
    const float4 dq = ((const float4*)a.d_rot)[i];
    const float2 cs = ((const float2*)a.cur_scores)[i];
    const float o_raw = a.opacity_raw[i];
    const float2 s_raw = ((const float2*)a.scaling_raw)[i];
    const float4 q = ((const float4*)a.rotation_raw)[i];
    // ---- backward of sigmoid / exp / normalize: the same expressions as sb_surfel_update_kernel (surfel_update.cu) ----
    const float y = 1.0f / (1.0f + expf(-o_raw));
    g[6] = g[6] * (1.0f - y) * y;
    g[7] = ds.x * expf(s_raw.x);
    g[8] = ds.y * expf(s_raw.y);
    const float len = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    if (len > 1e-12f) {
        const float4 r = make_float4(q.x / len, q.y / len, q.z / len, q.w / len);
        const float rd = r.x * dq.x + r.y * dq.y + r.z * dq.z + r.w * dq.w;
        g[9] = (dq.x - r.x * rd) / len; g[10] = (dq.y - r.y * rd) / len;
        g[11] = (dq.z - r.z * rd) / len; g[12] = (dq.w - r.w * rd) / len;
    } else {
        g[9] = dq.x / 1e-12f; g[10] = dq.y / 1e-12f; g[11] = dq.z / 1e-12f; g[12] = dq.w / 1e-12f;
    }
    float4* row = (float4*)(a.stage + iz * SB_DP_STAGE_F);
    row[0] = make_float4(g[0], g[1], g[2], g[3]);
    row[1] = make_float4(g[4], g[5], g[6], g[7]);
    row[2] = make_float4(g[8], g[9], g[10], g[11]);
    row[3] = make_float4(g[12], cs.x, cs.y, 0.0f);
}

struct SbDpReduceArgs {
    const float* peer_stage[SB_DP_MAX_RANKS];
    const uint8_t* peer_vis[SB_DP_MAX_RANKS];
    const float* global_scores;   // [P,2] local replica (identical on every rank)
    float* out;                   // [slice, SB_DP_OUT_F] this rank's result rows
    uint8_t* upd;                 // [slice] 0 = no view saw the surfel, else 1 + rank of the first arg-max of score 0
    uint8_t* peer_chunk[SB_DP_MAX_RANKS];   // every rank's chunk-touched bytes: the owner PUSHES its verdict to all of them
    int P, rank, n_ranks;         // ownership is cyclic over 256-surfel chunks: chunk c belongs to rank c % n_ranks
    SbDpSync sync;
};

// Rows that live on another GPU are fetched the way NVLink likes them: the CTA's 256 consecutive rows are one
// contiguous 16 KB span (stage and result rows are both 64 B), loaded with fully coalesced 512-byte warp requests into shared memory
// and only then picked apart per thread.  (Per-thread row loads -- 16-byte pieces at a 64-byte stride, 80 before the result row was packed -- ran the
// 2-GPU post kernel at ~110 GB/s of NVLink reads: 0.26 ms for 29 MB.)
__device__ __forceinline__ void sb_dp_reduce_body(const SbDpReduceArgs& a) {
    __shared__ float4 s_rows[2][256 * (SB_DP_STAGE_F / 4)];   // double buffered: the next rank's span arrives while this one is folded
    __shared__ uint32_t s_any;                                // bit q: rank q's view saw one of this CTA's surfels
    // CTA b of rank r folds chunk b * n_ranks + r (cyclic ownership: the surfels a window sees are clustered in index,
    // contiguous slices left the owners of the old end of the cloud idle and the others with 3x the average work)
    const int i0 = (blockIdx.x * a.n_ranks + a.rank) * 256;
    const int i = i0 + threadIdx.x;
    const bool mine = i < a.P;
    const size_t iz = (size_t)(mine ? i : 0);
    const int rows = max(0, min(256, a.P - i0));         // rows of this chunk that exist
    // every rank's visibility byte first: n_ranks independent peer loads in flight, not a chain of them
    uint32_t vis_bits = 0;
    if (threadIdx.x == 0) s_any = 0;
    __syncthreads();
    if (mine) {
#pragma unroll 4
        for (int q = 0; q < a.n_ranks; q++) vis_bits |= (sb_ld_peer_u8(a.peer_vis[q] + iz) != 0 ? 1u : 0u) << q;
    }
    const uint32_t warp_bits = __reduce_or_sync(0xffffffffu, vis_bits);
    if ((threadIdx.x & 31) == 0 && warp_bits) atomicOr(&s_any, warp_bits);
    const float glob0 = mine ? a.global_scores[2 * iz] : 0.0f;
    __syncthreads();
    uint32_t todo = s_any;
    // tell every rank whether anything in this chunk was seen: their post kernels then fetch the result span only if
    // so, without a peer round trip for the answer (remote stores, released by this kernel's hand-off signal)
    if ((int)threadIdx.x < a.n_ranks && rows > 0) a.peer_chunk[threadIdx.x][i0 >> 8] = todo ? 1 : 0;
    float acc[13];
#pragma unroll
    for (int c = 0; c < 13; c++) acc[c] = 0.0f;
    float s0_sum = 0.0f, s1_max = 0.0f, s0_max = -1.0f;
    int arg_rank = 0;
    auto fetch = [&](int q, int buf) {   // the CTA's 16-KB span of rank q's stage rows, coalesced 512-byte requests
        const float4* src = (const float4*)(a.peer_stage[q] + (size_t)i0 * SB_DP_STAGE_F);
        for (int k = threadIdx.x; k < rows * (SB_DP_STAGE_F / 4); k += 256) s_rows[buf][k] = sb_ld_peer_f4(src + k);
    };
    int buf = 0;
    if (todo) fetch(__ffs(todo) - 1, 0);
    while (todo) {   // ascending q = rank order = the order a single process would take the views in
        const int q = __ffs(todo) - 1;
        todo &= todo - 1;
        __syncthreads();                                  // span q has landed (and the other buffer is free again)
        if (todo) fetch(__ffs(todo) - 1, buf ^ 1);
        if ((vis_bits >> q) & 1u) {
            const float4* row = s_rows[buf] + threadIdx.x * (SB_DP_STAGE_F / 4);
            const float4 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];
            const float c0 = r3.y, c1 = r3.z;
            // weighting_grad with this view's score counted into the global score (add_records runs first), as in the
            // single-GPU pass; the global score itself is updated once, from the sum, by sb_dp_post_kernel
            const float gw = c0 / (fminf(fmaxf(glob0 + c0, 0.0f), 1e4f) + 1e-6f + c0);
            const float g[13] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x};
#pragma unroll
            for (int c = 0; c < 13; c++) acc[c] = __fadd_rn(acc[c], __fmul_rn(g[c], gw));   // a plain sum of the views' weighted rows
            s0_sum += c0;
            s1_max = fmaxf(s1_max, c1);
            if (c0 > s0_max) { s0_max = c0; arg_rank = q; }   // strict: the first view holding the maximum keeps it
        }
        buf ^= 1;
    }
    if (!mine) return;
    const int li = blockIdx.x * 256 + threadIdx.x;       // the chunk's place in the owner's result buffer
    a.upd[li] = vis_bits ? (uint8_t)(1 + arg_rank) : 0;
    if (!vis_bits) return;
    float4* o = (float4*)(a.out + (size_t)li * SB_DP_OUT_F);
    o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    o[2] = make_float4(acc[8], acc[9], acc[10], acc[11]);
    o[3] = make_float4(acc[12], s0_sum, s1_max, s0_max);
}

struct SbDpPostArgs {
    const float* peer_out[SB_DP_MAX_RANKS];
    const uint8_t* peer_upd[SB_DP_MAX_RANKS];
    float* param[5]; float* m[5]; float* v[5];
    const uint8_t* stable;
    float* local_scores; float* global_scores; float* kf_max; long long* kf_id;
    float lr[5];
    float b1, b2, eps;
    long long kf_ids[SB_DP_MAX_RANKS];   // keyframe id of every rank's view of this step
    const uint8_t* chunk;                // local chunk-touched bytes (written by the owners' reduce kernels)
    int P, n_ranks;
    SbDpSync sync;
};

#define SB_GRP(c) ((c) < 3 ? 0 : (c) < 6 ? 1 : (c) < 7 ? 2 : (c) < 9 ? 3 : 4)
#define SB_GRP_NC(g) ((g) == 0 ? 3 : (g) == 1 ? 3 : (g) == 2 ? 1 : (g) == 3 ? 2 : 4)
#define SB_GRP_C0(g) ((g) == 0 ? 0 : (g) == 1 ? 3 : (g) == 2 ? 6 : (g) == 3 ? 7 : 9)

__device__ __forceinline__ void sb_dp_post_body(const SbDpPostArgs& a) {
    __shared__ float4 s_rows[256 * (SB_DP_OUT_F / 4)];
    const int i0 = blockIdx.x * blockDim.x;              // one 256-surfel chunk: one owner
    const int i = i0 + threadIdx.x;
    const bool inside = i < a.P;
    const size_t iz = (size_t)(inside ? i : 0);
    const int chunk = blockIdx.x, owner = chunk % a.n_ranks, l0 = (chunk / a.n_ranks) * 256;
    // nobody saw any surfel of this chunk (the owner said so, into local memory): nothing to fetch, nothing to do
    if (a.chunk[chunk] == 0) return;
    // the owner's touched bytes and its span of result rows are requested together (one peer round trip, not two)
    const uint32_t ub = inside ? sb_ld_peer_u8(a.peer_upd[owner] + l0 + threadIdx.x) : 0u;
    {
        const int rows = min(256, a.P - i0);
        const float4* src = (const float4*)(a.peer_out[owner] + (size_t)l0 * SB_DP_OUT_F);
        for (int k = threadIdx.x; k < rows * (SB_DP_OUT_F / 4); k += 256) s_rows[k] = sb_ld_peer_f4(src + k);
    }
    __syncthreads();
    // no view saw the surfel: add_records with zero scores is the identity and Adam is masked out
    if (ub == 0) return;
    const float4* o = s_rows + threadIdx.x * (SB_DP_OUT_F / 4);
    const float4 r0 = o[0], r1 = o[1], r2 = o[2], r3 = o[3];
    const float g[13] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w, r3.x};
    const float s0_sum = r3.y, s1_max = r3.z, s0_max = r3.w;
    const long long kf_arg = a.kf_ids[ub - 1];   // the first view (in rank order) holding the score maximum
    const bool adam = !a.stable[i];

    // ---- every load before the first store (see sb_surfel_update_kernel) ----
    const float glob0 = a.global_scores[2 * iz], glob1 = a.global_scores[2 * iz + 1];
    const float loc0 = a.local_scores[2 * iz], loc1 = a.local_scores[2 * iz + 1], kfm = a.kf_max[i];
    float pv[13], mv[13], vv[13];
    if (adam) {
#pragma unroll
        for (int c = 0; c < 7; c++) {
            const int grp = SB_GRP(c), nc = SB_GRP_NC(grp);
            const size_t e = iz * nc + (c - SB_GRP_C0(grp));
            pv[c] = a.param[grp][e]; mv[c] = a.m[grp][e]; vv[c] = a.v[grp][e];
        }
        const float2 s2 = ((const float2*)a.param[3])[i], ms2 = ((const float2*)a.m[3])[i], vs2 = ((const float2*)a.v[3])[i];
        const float4 q4 = ((const float4*)a.param[4])[i], mq4 = ((const float4*)a.m[4])[i], vq4 = ((const float4*)a.v[4])[i];
        pv[7] = s2.x; pv[8] = s2.y; mv[7] = ms2.x; mv[8] = ms2.y; vv[7] = vs2.x; vv[8] = vs2.y;
        pv[9] = q4.x; pv[10] = q4.y; pv[11] = q4.z; pv[12] = q4.w;
        mv[9] = mq4.x; mv[10] = mq4.y; mv[11] = mq4.z; mv[12] = mq4.w;
        vv[9] = vq4.x; vv[10] = vq4.y; vv[11] = vq4.z; vv[12] = vq4.w;
    }

    // ---- add_records of the step's views, folded (gaussian_model.py:163-168, gaussian_base.py:388-392) ----
    a.local_scores[2 * iz] = loc0 + s0_sum;
    if (s1_max > loc1) a.local_scores[2 * iz + 1] = s1_max;
    a.global_scores[2 * iz] = fminf(fmaxf(glob0 + s0_sum, 0.0f), 1e4f);
    a.global_scores[2 * iz + 1] = fminf(fmaxf(glob1, 0.0f), 1e4f);
    if (kfm < s0_max) {
        a.kf_max[i] = s0_max;
        a.kf_id[i] = kf_arg;
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

__global__ void __launch_bounds__(256) sb_dp_stage_kernel(SbDpStageArgs a) {
    sb_dp_stage_body(a);
    sb_dp_signal(a.sync, 0);
}
__global__ void __launch_bounds__(256) sb_dp_reduce_kernel(SbDpReduceArgs a) {
    sb_dp_wait(a.sync, 0);
    sb_dp_reduce_body(a);
    sb_dp_signal(a.sync, 1);
}
__global__ void __launch_bounds__(256, 4) sb_dp_post_kernel(SbDpPostArgs a) {
    sb_dp_wait(a.sync, 1);
    sb_dp_post_body(a);
}

// ---- host side: one exchange handle per rank (layout of the peer-mapped buffer, peer pointers, step counter) ----
struct SbDpLayout { size_t stage, vis, out, upd, chunk, flags, counter, error, wait, total; int64_t slice_cap; };
static SbDpLayout sb_dp_layout(int64_t capacity, int n_ranks) {
    SbDpLayout L;
    const size_t cap = (size_t)(capacity > 0 ? capacity : 0);
    L.slice_cap = (int64_t)(((cap + 255) / 256 + n_ranks - 1) / n_ranks * 256);   // chunks per owner, rounded up
    size_t o = 0;
    L.stage = o;   o = sb_align(o + cap * SB_DP_STAGE_F * 4);
    L.vis = o;     o = sb_align(o + cap);
    L.out = o;     o = sb_align(o + (size_t)L.slice_cap * SB_DP_OUT_F * 4);
    L.upd = o;     o = sb_align(o + (size_t)L.slice_cap);
    L.chunk = o;   o = sb_align(o + (cap + 255) / 256);   // one byte per 256-surfel chunk: some view saw one of its surfels
    L.flags = o;   o = sb_align(o + 2 * SB_DP_MAX_RANKS * 4);
    L.counter = o; o = sb_align(o + 2 * 4);
    L.error = o;   o = sb_align(o + 4);
    L.wait = o;    o = sb_align(o + 2 * 8);
    L.total = o;
    return L;
}

struct SbDpHandle {
    int rank, n_ranks;
    int64_t capacity;
    SbDpLayout L;
    char* base[SB_DP_MAX_RANKS];
    uint32_t step;
};

static SbDpSync sb_dp_sync_of(const SbDpHandle* h) {
    SbDpSync y;
    memset(&y, 0, sizeof(y));
    for (int q = 0; q < h->n_ranks; q++) y.peer_flags[q] = (uint32_t*)(h->base[q] + h->L.flags);
    y.counter = (uint32_t*)(h->base[h->rank] + h->L.counter);
    y.error = (uint32_t*)(h->base[h->rank] + h->L.error);
    y.wait_ns = (unsigned long long*)(h->base[h->rank] + h->L.wait);
    y.step = h->step;
    y.rank = h->rank;
    y.n_ranks = h->n_ranks;
    return y;
}

// used by sb200_backward_update(mode 1) in surfel_update.cu
int sb_dp_begin_step(void* handle, int P, float** stage, uint8_t** vis, SbDpSync* sync) {
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h) { sb_set_error("multi-GPU exchange handle is NULL (sb200_dp_create)"); return 1; }
    if (P > h->capacity) { sb_set_error("exchange: %d surfels exceed the capacity %lld", P, (long long)h->capacity); return 1; }
    h->step += 1;   // every rank calls in lockstep: the same number everywhere
    *stage = (float*)(h->base[h->rank] + h->L.stage);
    *vis = (uint8_t*)(h->base[h->rank] + h->L.vis);
    *sync = sb_dp_sync_of(h);
    return 0;
}

extern "C" {

size_t sb200_dp_buffer_bytes(int64_t capacity, int n_ranks) {
    if (n_ranks < 1 || n_ranks > SB_DP_MAX_RANKS) return 0;
    return sb_dp_layout(capacity, n_ranks).total;
}

void* sb200_dp_create(int rank, int n_ranks, int64_t capacity, const void* const* peer_bases) {
    if (n_ranks < 1 || n_ranks > SB_DP_MAX_RANKS || rank < 0 || rank >= n_ranks || capacity < 0 || !peer_bases) {
        sb_set_error("dp_create: bad rank / rank count / capacity");
        return nullptr;
    }
    SbDpHandle* h = new SbDpHandle();
    h->rank = rank; h->n_ranks = n_ranks; h->capacity = capacity; h->step = 0;
    h->L = sb_dp_layout(capacity, n_ranks);
    for (int q = 0; q < n_ranks; q++) {
        if (!peer_bases[q]) { sb_set_error("dp_create: NULL buffer of rank %d", q); delete h; return nullptr; }
        h->base[q] = (char*)peer_bases[q];
    }
    return h;
}

void sb200_dp_destroy(void* handle) { delete (SbDpHandle*)handle; }

int sb200_dp_slice(void* handle, int P) {   // rows of the result buffer a rank needs for P surfels
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h) return 0;
    const int chunks = (P + 255) / 256;
    return (chunks + h->n_ranks - 1) / h->n_ranks * 256;
}

int sb200_dp_error(void* handle, void* stream) {
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h) return 1;
    uint32_t e = 0;
    if (cudaMemcpyAsync(&e, h->base[h->rank] + h->L.error, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream) != cudaSuccess ||
        cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) { sb_set_error("dp_error: read failed"); return 2; }
    if (e) { sb_set_error("multi-GPU exchange: a rank waited 20 s for a peer that never signalled"); return 3; }
    return 0;
}

int sb200_dp_wait_ns(void* handle, uint64_t out[2], void* stream) {
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h || !out) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaMemcpyAsync(out, h->base[h->rank] + h->L.wait, 16, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
        cudaMemsetAsync(h->base[h->rank] + h->L.wait, 0, 16, s) != cudaSuccess ||
        cudaStreamSynchronize(s) != cudaSuccess) { sb_set_error("dp_wait_ns: read failed"); return 2; }
    return 0;
}

int sb200_dp_stage(void* handle, int P, const float* d_means3D, const float* d_colors, const float* d_opacity,
                   const float* d_scales, const float* d_rotations, const float* cur_scores, const int* radii,
                   const float* opacity_raw, const float* scaling_raw, const float* rotation_raw, void* stream) {
    if (P < 0) { sb_set_error("dp_stage: bad P"); return 1; }
    if (P > 0 && (!d_means3D || !d_colors || !d_opacity || !d_scales || !d_rotations || !cur_scores || !radii ||
                  !opacity_raw || !scaling_raw || !rotation_raw)) { sb_set_error("dp_stage: NULL input"); return 1; }
    if ((((uintptr_t)d_scales | (uintptr_t)cur_scores | (uintptr_t)scaling_raw) & 7) ||
        (((uintptr_t)d_rotations | (uintptr_t)rotation_raw) & 15)) {
        sb_set_error("dp_stage: [P,2] arrays must be 8-byte, [P,4] arrays 16-byte aligned");
        return 1;
    }
    SbDpStageArgs a = {d_means3D, d_colors, d_opacity, d_scales, d_rotations, cur_scores, radii,
                       opacity_raw, scaling_raw, rotation_raw, nullptr, nullptr, P};
    if (sb_dp_begin_step(handle, P, &a.stage, &a.vis, &a.sync)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    SbStageScope sc(SB200_STAGE_UPDATE, s);
    sb_dp_stage_kernel<<<(P + 255) / 256 + (P == 0), 256, 0, s>>>(a);   // at least one CTA: it carries the signal
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int sb200_dp_reduce(void* handle, int P, const float* global_scores, void* stream) {
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h || P < 0 || P > h->capacity) { sb_set_error("dp_reduce: bad handle / P"); return 1; }
    if (P > 0 && !global_scores) { sb_set_error("dp_reduce: NULL input"); return 1; }
    SbDpReduceArgs a;
    memset(&a, 0, sizeof(a));
    for (int q = 0; q < h->n_ranks; q++) {
        a.peer_stage[q] = (const float*)(h->base[q] + h->L.stage);
        a.peer_vis[q] = (const uint8_t*)(h->base[q] + h->L.vis);
        a.peer_chunk[q] = (uint8_t*)(h->base[q] + h->L.chunk);
    }
    a.global_scores = global_scores;
    a.out = (float*)(h->base[h->rank] + h->L.out);
    a.upd = (uint8_t*)(h->base[h->rank] + h->L.upd);
    a.P = P; a.rank = h->rank; a.n_ranks = h->n_ranks;
    a.sync = sb_dp_sync_of(h);
    cudaStream_t s = (cudaStream_t)stream;
    SbStageScope sc(SB200_STAGE_DP_REDUCE, s);
    const int chunks = (P + 255) / 256;
    const int mine = chunks > h->rank ? (chunks - h->rank + h->n_ranks - 1) / h->n_ranks : 0;   // chunks r, r + N, ...
    sb_dp_reduce_kernel<<<mine > 0 ? mine : 1, 256, 0, s>>>(a);   // at least one CTA: it carries the hand-off
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int sb200_dp_post(void* handle, int P, const int64_t* kf_ids, float* const* params, float* const* exp_avg, float* const* exp_avg_sq,
                  const uint8_t* stable_mask, float* local_scores, float* global_scores, float* kf_max_scores,
                  int64_t* kf_id, const float* lr, float b1, float b2, float eps, void* stream) {
    SbDpHandle* h = (SbDpHandle*)handle;
    if (!h || P < 0 || P > h->capacity) { sb_set_error("dp_post: bad handle / P"); return 1; }
    if (!kf_ids || !params || !exp_avg || !exp_avg_sq || !lr || (P > 0 && (!stable_mask || !local_scores || !global_scores ||
                                                                 !kf_max_scores || !kf_id))) {
        sb_set_error("dp_post: NULL input");
        return 1;
    }
    if ((((uintptr_t)params[3] | (uintptr_t)exp_avg[3] | (uintptr_t)exp_avg_sq[3]) & 7) ||
        (((uintptr_t)params[4] | (uintptr_t)exp_avg[4] | (uintptr_t)exp_avg_sq[4]) & 15)) {
        sb_set_error("dp_post: scaling arrays must be 8-byte, rotation arrays 16-byte aligned");
        return 1;
    }
    SbDpPostArgs a;
    memset(&a, 0, sizeof(a));
    for (int q = 0; q < h->n_ranks; q++) {
        a.peer_out[q] = (const float*)(h->base[q] + h->L.out);
        a.peer_upd[q] = (const uint8_t*)(h->base[q] + h->L.upd);
        a.kf_ids[q] = kf_ids[q];
    }
    for (int g = 0; g < 5; g++) { a.param[g] = params[g]; a.m[g] = exp_avg[g]; a.v[g] = exp_avg_sq[g]; a.lr[g] = lr[g]; }
    a.stable = stable_mask; a.local_scores = local_scores; a.global_scores = global_scores; a.kf_max = kf_max_scores;
    a.kf_id = (long long*)kf_id; a.b1 = b1; a.b2 = b2; a.eps = eps; a.P = P; a.n_ranks = h->n_ranks;
    a.chunk = (const uint8_t*)(h->base[h->rank] + h->L.chunk);
    a.sync = sb_dp_sync_of(h);
    cudaStream_t s = (cudaStream_t)stream;
    SbStageScope sc(SB200_STAGE_DP_POST, s);
    sb_dp_post_kernel<<<(P + 255) / 256 + (P == 0), 256, 0, s>>>(a);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // extern "C"
