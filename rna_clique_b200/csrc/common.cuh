// common.cuh — context, error plumbing and device-buffer helpers shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/rcq.h"

#define RCQ_SMS 148  // B200: grids are sized in multiples of the SM count

extern thread_local std::string g_rcq_err;

static inline int rcq_fail(int code, const char* fmt, const char* a = "", const char* b = "") {
    char buf[512];
    snprintf(buf, sizeof buf, fmt, a, b);
    g_rcq_err = buf;
    return code;
}

#define CU(x)                                                                             \
    do {                                                                                  \
        cudaError_t e_ = (x);                                                             \
        if (e_ != cudaSuccess) {                                                          \
            char b_[512];                                                                 \
            snprintf(b_, sizeof b_, "%s:%d: %s: %s", __FILE__, __LINE__, #x,              \
                     cudaGetErrorString(e_));                                             \
            g_rcq_err = b_;                                                               \
            return RCQ_ERR_CUDA;                                                          \
        }                                                                                 \
    } while (0)
#define RC(x)                     \
    do {                          \
        int r_ = (x);             \
        if (r_ != 0) return r_;   \
    } while (0)

struct rcq_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaMemPool_t pool = nullptr;   // this context's own stream-ordered pool (no cross-stream reuse stalls)
    int64_t launches = 0;
    // pinned staging for small device->host result reads
    int64_t* h_pin = nullptr;  // 64 int64 (pinned, mapped: written by k_publish, no copy engine involved)
    int64_t* h_pin_dev = nullptr;  // device-side address of h_pin
    char* h_stage = nullptr;   // pinned, mapped staging for the small per-sample host arrays
    char* h_stage_dev = nullptr;
    size_t h_stage_cap = 0;
    int64_t* d_ctr = nullptr;  // 64 int64 device counters
    // debug: gapped HSPs of the last search (device, 12 int32 each)
    int32_t* dbg_hsp = nullptr;
    int64_t dbg_n = 0;
    int dbg_keep = 0;
    // transient arena (talloc/arena_reset below): scratch of one API call, bump-allocated
    char* arena = nullptr;
    size_t arena_cap = 0, arena_used = 0, arena_need = 0;
    std::vector<void*> arena_overflow;
    // fallback rows of the gapped kernel (search.cu: scratch_pool)
    int32_t* gs_rows = nullptr;
    unsigned int* gs_locks = nullptr;
    int64_t gs_per_slot = 0;
    int32_t gs_slots = 0;
    int dust_on = 1;              // symmetric DUST soft-masking of seeds in samples built on this context (blastn default)
    int gapped_occ[3] = {0, 0, 0};   // resident blocks per SM of the gapped kernels (occupancy query, cached)
    // per-stage device time (CUDA events on this stream), accumulated across calls:
    // 0 pack+index, 1 seed join, 2 ungapped, 3 sort/group, 4 gapped DP, 5 selection+table, 6 phase 2
    double stage_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t stage_calls[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaEvent_t ev[16] = {};
    int pending[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// stage timers: tic/toc record events; collect() folds every pending pair after a stream sync
static inline void stage_tic(rcq_ctx* c, int st) { cudaEventRecord(c->ev[2 * st], c->stream); }
static inline void stage_toc(rcq_ctx* c, int st) { cudaEventRecord(c->ev[2 * st + 1], c->stream); c->pending[st] = 1; }
static inline void stage_collect(rcq_ctx* c) {
    for (int st = 0; st < 8; st++)
        if (c->pending[st]) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, c->ev[2 * st], c->ev[2 * st + 1]) == cudaSuccess) {
                c->stage_ms[st] += ms;
                c->stage_calls[st]++;
            }
            c->pending[st] = 0;
        }
}

// stream-ordered device allocation (pool keeps memory; no OS round trips in steady state)
template <typename T>
static inline int dalloc(rcq_ctx* c, T** p, int64_t n) {
    if (n < 1) n = 1;
    CU(cudaMallocFromPoolAsync((void**)p, (size_t)n * sizeof(T), c->pool, c->stream));
    return 0;
}
template <typename T>
static inline void dfree(rcq_ctx* c, T* p) {
    if (p) cudaFreeAsync((void*)p, c->stream);
}

// Transient scratch of one API call: bump allocation out of one per-context slab, so the steady
// state makes no allocator calls at all (a pool that has to grow mid-step maps new physical memory,
// which stalls every stream of the device for tens to hundreds of milliseconds).  Every entry point
// that uses talloc starts with arena_reset and ends with a stream synchronise.  A request that does
// not fit falls back to the pool for this call and the slab is regrown at the next reset.
static inline int arena_reset(rcq_ctx* c) {
    for (void* p : c->arena_overflow) cudaFreeAsync(p, c->stream);
    c->arena_overflow.clear();
    if (c->arena_need > c->arena_cap) {
        CU(cudaStreamSynchronize(c->stream));
        if (c->arena) CU(cudaFree(c->arena));
        c->arena = nullptr;
        c->arena_cap = c->arena_need + c->arena_need / 4 + (64u << 20);
        CU(cudaMalloc((void**)&c->arena, c->arena_cap));
    }
    c->arena_used = 0;
    c->arena_need = 0;
    return 0;
}
template <typename T>
static inline int talloc(rcq_ctx* c, T** p, int64_t n) {
    if (n < 1) n = 1;
    const size_t bytes = ((size_t)n * sizeof(T) + 255) & ~(size_t)255;
    c->arena_need += bytes;
    if (c->arena_used + bytes <= c->arena_cap) {
        *p = (T*)(c->arena + c->arena_used);
        c->arena_used += bytes;
        return 0;
    }
    CU(cudaMallocFromPoolAsync((void**)p, bytes, c->pool, c->stream));
    c->arena_overflow.push_back((void*)*p);
    return 0;
}
template <typename T>
static inline void tfree(rcq_ctx*, T*) {}   // released wholesale by the next arena_reset

// Small host<->device traffic bypasses the copy engines (a bulk host->device copy in flight would make
// every small cudaMemcpyAsync wait its turn behind it): counters are published by a one-block kernel
// into mapped pinned memory, small host arrays are read by kernels from mapped pinned staging.
static __global__ void k_publish(const int64_t* __restrict__ src, volatile int64_t* dst, int n) {
    if ((int)threadIdx.x < n) dst[threadIdx.x] = src[threadIdx.x];
    __threadfence_system();
}
static inline int publish_counters(rcq_ctx* c, const int64_t* d_src, int n) {
    k_publish<<<1, 64, 0, c->stream>>>(d_src, c->h_pin_dev, n);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}
static inline int stage_reserve(rcq_ctx* c, size_t bytes) {
    if (bytes <= c->h_stage_cap) return 0;
    CU(cudaStreamSynchronize(c->stream));
    if (c->h_stage) CU(cudaFreeHost(c->h_stage));
    c->h_stage = nullptr; c->h_stage_cap = 0;
    const size_t cap = bytes + bytes / 2 + (1u << 20);
    CU(cudaHostAlloc((void**)&c->h_stage, cap, cudaHostAllocMapped));
    CU(cudaHostGetDevicePointer((void**)&c->h_stage_dev, c->h_stage, 0));
    c->h_stage_cap = cap;
    return 0;
}

#define LAUNCH(ctx, kern, grid, block, smem, ...)                       \
    do {                                                                \
        kern<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);  \
        (ctx)->launches++;                                              \
        CU(cudaGetLastError());                                         \
    } while (0)

static inline int grid_for(int64_t n, int block, int per_thread = 1) {
    int64_t g = (n + (int64_t)block * per_thread - 1) / ((int64_t)block * per_thread);
    if (g < 1) g = 1;
    if (g > (1 << 30)) g = 1 << 30;
    return (int)g;
}

// ---- scoring constants (blastn megablast defaults) -------------------
// Scores are kept doubled inside the gapped kernel so the linear gap cost 2.5 is integral.
#define RCQ_WORD 28          // contiguous exact seed length
#define RCQ_KMER 16          // minimizer k-mer (any 28-mer window holds 13 of them)
#define RCQ_WIN 13           // RCQ_WORD - RCQ_KMER + 1
#define RCQ_XDROP_UNGAPPED 10   // 20 bits  -> floor(20 ln2 / 1.28)
#define RCQ_XDROP_GAPPED2 108   // 100 bits -> floor(100 ln2 / 1.28) = 54, doubled
#define RCQ_MATCH2 2
#define RCQ_MISMATCH2 (-4)
#define RCQ_GAP2 (-5)
#define RCQ_MAX_GAPPED_PER_GROUP 32
#define RCQ_KA_LAMBDA 1.28
#define RCQ_KA_K 0.46
#define RCQ_KA_ALPHA 1.5
#define RCQ_KA_BETA (-2.0)

// invertible 32-bit mixer (so equal hashes <=> equal k-mers): murmur3 finaliser
__host__ __device__ static inline uint32_t rcq_mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x85ebca6bu; x ^= x >> 13; x *= 0xc2b2ae35u; x ^= x >> 16;
    return x;
}
// index sort key of a minimizer hash (bijective, so equal keys <=> equal canonical 16-mers)
__host__ __device__ static inline uint32_t rcq_index_key(uint32_t h) { return rcq_mix32(h ^ 0x9E3779B9u); }
