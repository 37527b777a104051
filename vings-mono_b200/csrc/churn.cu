// N3: surfel add / prune churn on the GPU.
//
//   sb_knn_dist2_kernel      replaces distCUDA2 (scripts/gaussian/gaussian_utils.py:12-16): the reference copies the
//                            <= 40 000 new points to the host, builds a scipy KD-tree, queries k=4 and copies the mean
//                            of the squared distances to the 3 nearest neighbours back -- a GPU<->CPU round trip and
//                            ~100 ms of host time per keyframe.  Here: tiled brute force in shared memory (N^2 = 1.6e9
//                            distance evaluations for 40 k points, ~0.3 ms), 4 smallest kept in registers; the point
//                            itself is one of the 4 (distance 0), exactly like KDTree.query(points, k=4)[:, 1:].
//   sb_prune_*_kernel        replaces prune_tensors_from_optimizer + update_records(mode="prune")
//                            (scripts/gaussian/gaussian_base.py:96-112): the reference boolean-indexes 15 tensors
//                            (5 parameters, 10 Adam moments) and 5 bookkeeping tensors one by one (~4 launches, one
//                            allocation and one host sync each).  Here: one stable compaction of every per-surfel array
//                            in three launches (block counts, scan of the counts, fused gather of all arrays) from one
//                            capacity buffer set into the other; one 4-byte read-back of the surviving count.
#include "common.cuh"
#include <mutex>
#include "../../include/surfel_b200.h"

#define KNN_THREADS 256
#define KNN_TILE 1024

__device__ __forceinline__ void knn_insert(float d, float& b0, float& b1, float& b2, float& b3) {
    if (d < b3) {
        if (d < b2) {
            b3 = b2;
            if (d < b1) {
                b2 = b1;
                if (d < b0) { b1 = b0; b0 = d; } else { b1 = d; }
            } else {
                b2 = d;
            }
        } else {
            b3 = d;
        }
    }
}

__global__ void __launch_bounds__(KNN_THREADS)
sb_knn_dist2_kernel(int N, const float* __restrict__ pts, float* __restrict__ out) {
    __shared__ float sx[KNN_TILE], sy[KNN_TILE], sz[KNN_TILE];
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (i < N) { qx = pts[3 * (size_t)i]; qy = pts[3 * (size_t)i + 1]; qz = pts[3 * (size_t)i + 2]; }
    const float INF = __int_as_float(0x7f800000);
    float b0 = INF, b1 = INF, b2 = INF, b3 = INF;
    for (int t0 = 0; t0 < N; t0 += KNN_TILE) {
        const int cnt = min(KNN_TILE, N - t0);
        __syncthreads();
        for (int e = threadIdx.x; e < cnt; e += KNN_THREADS) {
            sx[e] = pts[3 * (size_t)(t0 + e)];
            sy[e] = pts[3 * (size_t)(t0 + e) + 1];
            sz[e] = pts[3 * (size_t)(t0 + e) + 2];
        }
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const float dx = sx[j] - qx, dy = sy[j] - qy, dz = sz[j] - qz;
            const float d = dx * dx + dy * dy + dz * dz;
            knn_insert(d, b0, b1, b2, b3);
        }
    }
    // b0 is the point itself (or a duplicate at distance 0): KDTree.query(k=4)[:, 1:]
    if (i < N) out[i] = (b1 + b2 + b3) / 3.0f;
}

// ---- stable compaction of every per-surfel array ----
#define PR_THREADS 256
#define PR_ROWS 1024  // rows per CTA (4 per thread)

struct SbPruneArrays {
    const uint8_t* src[SB200_PRUNE_MAX_ARRAYS];
    uint8_t* dst[SB200_PRUNE_MAX_ARRAYS];
    int bytes[SB200_PRUNE_MAX_ARRAYS];  // bytes per row (word copies when a multiple of 4: buffers are 256-B aligned)
    int n;
};

__global__ void __launch_bounds__(PR_THREADS)
sb_prune_count_kernel(int P, const uint8_t* __restrict__ del, uint32_t* __restrict__ block_count) {
    __shared__ uint32_t s_w[PR_THREADS / 32];
    const int base = blockIdx.x * PR_ROWS;
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < PR_ROWS / PR_THREADS; k++) {
        const int r = base + k * PR_THREADS + threadIdx.x;
        c += (r < P && !del[r]);
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < PR_THREADS / 32; w++) t += s_w[w];
        block_count[blockIdx.x] = t;
    }
}

// one CTA: exclusive scan of the block counts in place; total -> block_count[nblocks]
__global__ void __launch_bounds__(1024) sb_prune_scan_kernel(int nblocks, uint32_t* __restrict__ block_count) {
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nblocks; b0 += 1024) {
        const int b = b0 + tid;
        const uint32_t v = b < nblocks ? block_count[b] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_w[wid] = incl;
        __syncthreads();
        uint32_t add = s_carry;
        for (int w = 0; w < wid; w++) add += s_w[w];
        if (b < nblocks) block_count[b] = add + incl - v;
        __syncthreads();
        if (tid == 1023) s_carry = add + incl;
        __syncthreads();
    }
    if (tid == 0) block_count[nblocks] = s_carry;
}

__global__ void __launch_bounds__(PR_THREADS)
sb_prune_gather_kernel(int P, const uint8_t* __restrict__ del, const uint32_t* __restrict__ block_base,
                       SbPruneArrays A) {
    __shared__ uint32_t s_w[PR_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // thread t owns the 4 consecutive rows base + 4t .. 4t+3, so the order of the rows is the order of the threads
    const int r0 = blockIdx.x * PR_ROWS + tid * (PR_ROWS / PR_THREADS);
    bool keep[PR_ROWS / PR_THREADS];
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < PR_ROWS / PR_THREADS; k++) {
        keep[k] = (r0 + k) < P && !del[r0 + k];
        c += keep[k];
    }
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[wid] = incl;
    __syncthreads();
    uint32_t d = block_base[blockIdx.x] + incl - c;
    for (int w = 0; w < wid; w++) d += s_w[w];
#pragma unroll
    for (int k = 0; k < PR_ROWS / PR_THREADS; k++) {
        if (!keep[k]) continue;
        const size_t r = (size_t)(r0 + k);
        for (int a = 0; a < A.n; a++) {
            const int nb = A.bytes[a];
            const uint8_t* s = A.src[a] + r * nb;
            uint8_t* t = A.dst[a] + (size_t)d * nb;
            if ((nb & 3) == 0) {
                for (int x = 0; x < nb / 4; x++) ((uint32_t*)t)[x] = ((const uint32_t*)s)[x];
            } else {
                for (int x = 0; x < nb; x++) t[x] = s[x];
            }
        }
        d++;
    }
}

extern "C" {

int sb200_knn_mean_dist2(int N, const float* points, float* out, void* stream) {
    if (N < 0) { sb_set_error("knn: bad N"); return 1; }
    if (N == 0) return 0;
    if (!points || !out) { sb_set_error("knn: NULL input"); return 1; }
    SbStageScope sc(SB200_STAGE_CHURN, (cudaStream_t)stream);
    sb_knn_dist2_kernel<<<(N + KNN_THREADS - 1) / KNN_THREADS, KNN_THREADS, 0, (cudaStream_t)stream>>>(N, points, out);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

size_t sb200_prune_scratch_bytes(int64_t P) {
    const size_t nblocks = (size_t)((P > 0 ? P : 0) + PR_ROWS - 1) / PR_ROWS;
    return sb_align((nblocks + 2) * sizeof(uint32_t));
}

int sb200_prune_rows(int P, const uint8_t* delete_mask, int n_arrays, const void* const* src, void* const* dst,
                     const int* row_bytes, void* scratch, int* kept_host, void* stream) {
    if (P < 0 || n_arrays < 0 || n_arrays > SB200_PRUNE_MAX_ARRAYS) { sb_set_error("prune: bad P / n_arrays"); return 1; }
    if (!kept_host) { sb_set_error("prune: kept_host is NULL"); return 1; }
    *kept_host = 0;
    if (P == 0) return 0;
    if (!delete_mask || !scratch || (n_arrays > 0 && (!src || !dst || !row_bytes))) { sb_set_error("prune: NULL input"); return 1; }
    SbPruneArrays A;
    A.n = n_arrays;
    for (int a = 0; a < n_arrays; a++) {
        if (row_bytes[a] <= 0) { sb_set_error("prune: row_bytes[%d] must be positive", a); return 1; }
        if (row_bytes[a] % 4 == 0 && (((uintptr_t)src[a] | (uintptr_t)dst[a]) & 3)) { sb_set_error("prune: array %d is not 4-byte aligned", a); return 1; }
        if (!src[a] || !dst[a] || src[a] == dst[a]) { sb_set_error("prune: array %d needs distinct src and dst buffers", a); return 1; }
        A.src[a] = (const uint8_t*)src[a];
        A.dst[a] = (uint8_t*)dst[a];
        A.bytes[a] = row_bytes[a];
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int nblocks = (P + PR_ROWS - 1) / PR_ROWS;
    uint32_t* counts = (uint32_t*)scratch;
    {
        SbStageScope sc(SB200_STAGE_CHURN, s);
        sb_prune_count_kernel<<<nblocks, PR_THREADS, 0, s>>>(P, delete_mask, counts);
        sb_prune_scan_kernel<<<1, 1024, 0, s>>>(nblocks, counts);
        sb_prune_gather_kernel<<<nblocks, PR_THREADS, 0, s>>>(P, delete_mask, counts, A);
        sb_count_launch(3);
    }
    SB_CUDA_CHECK(cudaGetLastError());
    SB_CUDA_CHECK(cudaMemcpyAsync(kept_host, counts + nblocks, 4, cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaStreamSynchronize(s));
    return 0;
}

}  // extern "C"

// ---- sb200_prune_rows_static: the same compaction for callers that must not allocate -------------------------------
// No counterpart in the reference (include/surfel_b200.h).  The block
// counts live in a file-scope device array instead of caller scratch: 64 KiB + 8 B of every GPU that loads the library,
// held until exit.  Host code cannot use the symbol's name as a pointer (it is a host-side stand-in, not a device
// address): the address is taken once per device with cudaGetSymbolAddress, with that device current.
static __device__ __align__(256) uint32_t g_prune_static_counts[SB200_PRUNE_STATIC_MAX_ROWS / PR_ROWS + 2];

static int sb_prune_static_scratch(uint32_t** out) {
    static std::mutex mu;
    static uint32_t* ptr_dev[SB_MAX_DEVICES] = {};
    int dev = -1;
    SB_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= SB_MAX_DEVICES) { sb_set_error("prune_static: device %d out of range", dev); return 1; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr_dev[dev]) {
        uint32_t* p = nullptr;
        SB_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_prune_static_counts));
        if (!p) { sb_set_error("prune_static: cudaGetSymbolAddress returned NULL"); return 2; }
        ptr_dev[dev] = p;
    }
    *out = ptr_dev[dev];
    return 0;
}

extern "C" int sb200_prune_rows_static(int P, const uint8_t* delete_mask, int n_arrays, const void* const* src,
                                       void* const* dst, const int* row_bytes, int* kept_dev, void* stream) {
    if (P < 0 || n_arrays < 0 || n_arrays > SB200_PRUNE_MAX_ARRAYS) { sb_set_error("prune_static: bad P / n_arrays"); return 1; }
    if (P > SB200_PRUNE_STATIC_MAX_ROWS) {
        sb_set_error("prune_static: P = %d exceeds SB200_PRUNE_STATIC_MAX_ROWS = %d (the static scratch is fixed; use "
                     "sb200_prune_rows with caller scratch)", P, (int)SB200_PRUNE_STATIC_MAX_ROWS);
        return 1;
    }
    if (!kept_dev) { sb_set_error("prune_static: kept_dev is NULL"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    if (P == 0) {
        SB_CUDA_CHECK(cudaMemsetAsync(kept_dev, 0, 4, s));
        return 0;
    }
    if (!delete_mask || (n_arrays > 0 && (!src || !dst || !row_bytes))) { sb_set_error("prune_static: NULL input"); return 1; }
    SbPruneArrays A;
    A.n = n_arrays;
    for (int a = 0; a < n_arrays; a++) {
        if (row_bytes[a] <= 0) { sb_set_error("prune_static: row_bytes[%d] must be positive", a); return 1; }
        if (!src[a] || !dst[a] || src[a] == dst[a]) { sb_set_error("prune_static: array %d needs distinct src and dst buffers", a); return 1; }
        if (row_bytes[a] % 4 == 0 && (((uintptr_t)src[a] | (uintptr_t)dst[a]) & 3)) { sb_set_error("prune_static: array %d is not 4-byte aligned", a); return 1; }
        A.src[a] = (const uint8_t*)src[a];
        A.dst[a] = (uint8_t*)dst[a];
        A.bytes[a] = row_bytes[a];
    }
    uint32_t* counts = nullptr;
    if (int rc = sb_prune_static_scratch(&counts)) return rc;
    const int nblocks = (P + PR_ROWS - 1) / PR_ROWS;
    // Nothing in the array has to start at zero: the count kernel writes counts[0 .. nblocks) and the scan writes
    // counts[nblocks] before anything reads them, so the previous call's contents are dead.
    sb_prune_count_kernel<<<nblocks, PR_THREADS, 0, s>>>(P, delete_mask, counts);
    sb_prune_scan_kernel<<<1, 1024, 0, s>>>(nblocks, counts);
    sb_prune_gather_kernel<<<nblocks, PR_THREADS, 0, s>>>(P, delete_mask, counts, A);
    sb_count_launch(3);
    SB_CUDA_CHECK(cudaGetLastError());
    SB_CUDA_CHECK(cudaMemcpyAsync(kept_dev, counts + nblocks, 4, cudaMemcpyDeviceToDevice, s));
    return 0;
}
