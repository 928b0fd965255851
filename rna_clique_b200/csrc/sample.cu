// sample.cu — kernels (1) and (2): 2-bit packing of a sample's top-gene transcripts and the
// sorted minimizer seed index.  Replaces `makeblastdb` per sample
// (reference find_all_pairs.py:120-159).  HBM-bound streaming kernels; algorithmic bytes:
//   pack   : 1 B/nt read + 0.25 B/nt written (+0.25 rc +2x1/32 masks)
//   index  : 0.25 B/nt read, 8 B per selected minimizer written, then an 8-B-key radix sort.
#include <cub/cub.cuh>
#include "sample.cuh"

#define GUARD_FRONT 64
#define GUARD_BACK 128

__device__ __forceinline__ uint32_t base_code(uint32_t c) {  // 0..3, or 4 = not ACGT(U)
    uint32_t u = c & 0xDFu;
    return u == 'A' ? 0u : u == 'C' ? 1u : u == 'G' ? 2u : (u == 'T' || u == 'U') ? 3u : 4u;
}

// Warp per transcript.  Each lane turns 16 letters into one 2-bit word per iteration; the 16
// letters come from two aligned 16-byte vector loads funnel-shifted to the transcript's byte
// alignment (coalesced: a warp iteration reads one contiguous 512(+16)-byte span).
__global__ void k_pack(const uint8_t* __restrict__ ascii, int64_t ascii_len,
                       const int64_t* __restrict__ starts, const int32_t* __restrict__ aoff,
                       const int32_t* __restrict__ len, int32_t n_seq, uint32_t* __restrict__ fwd, uint32_t* __restrict__ mask,
                       unsigned long long* __restrict__ n_masked) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < n_seq; t += nwarps) {
        int64_t src = starts[t];
        int32_t n = len[t];
        int64_t dst = aoff[t];
        unsigned long long masked = 0;
        for (int32_t c = 0; c < n; c += 512) {
            int32_t b0 = c + lane * 16;
            uint32_t word = 0, mbits = 0;
            if (b0 < n) {
                int64_t a = src + b0;
                uint64_t A, B;  // the 16 letters, little-endian
                int64_t al = a & ~15ll;
                if (al + 32 <= ascii_len && (((uintptr_t)ascii) & 15) == 0) {
                    uint4 v0 = __ldg((const uint4*)(ascii + al));
                    uint4 v1 = __ldg((const uint4*)(ascii + al + 16));
                    uint64_t q0 = ((uint64_t)v0.y << 32) | v0.x, q1 = ((uint64_t)v0.w << 32) | v0.z;
                    uint64_t q2 = ((uint64_t)v1.y << 32) | v1.x, q3 = ((uint64_t)v1.w << 32) | v1.z;
                    int s = (int)(a - al);
                    if (s >= 8) { q0 = q1; q1 = q2; q2 = q3; s -= 8; }
                    if (s == 0) { A = q0; B = q1; }
                    else {
                        A = (q0 >> (8 * s)) | (q1 << (64 - 8 * s));
                        B = (q1 >> (8 * s)) | (q2 << (64 - 8 * s));
                    }
                } else {
                    A = 0; B = 0;
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        A |= (uint64_t)((a + k < ascii_len) ? ascii[a + k] : 'N') << (8 * k);
                        B |= (uint64_t)((a + 8 + k < ascii_len) ? ascii[a + 8 + k] : 'N') << (8 * k);
                    }
                }
                int lim = min(16, n - b0);
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    uint32_t chk = (uint32_t)((k < 8 ? A >> (8 * k) : B >> (8 * (k - 8))) & 0xFF);
                    uint32_t code = (k < lim) ? base_code(chk) : 4u;
                    if (code > 3u) {
                        mbits |= 1u << k;
                        if (k < lim) masked++;
                    } else {
                        word |= code << (2 * k);
                    }
                }
                fwd[(dst + b0) >> 4] = word;
            } else {
                mbits = 0xFFFFu;
            }
            uint32_t other = __shfl_down_sync(0xffffffffu, mbits, 1);
            if ((lane & 1) == 0 && b0 < n) mask[(dst + b0) >> 5] = mbits | (other << 16);
        }
        for (int o = 16; o; o >>= 1) masked += __shfl_down_sync(0xffffffffu, masked, o);
        if (lane == 0 && masked) atomicAdd(n_masked, masked);
    }
}

__device__ __forceinline__ uint32_t rev2_32(uint32_t x) {  // reverse the 16 2-bit groups
    x = __brev(x);
    return ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
}

// Per-transcript reverse complement of the packed arena (and of the mask).
__global__ void k_revcomp(const uint32_t* __restrict__ fwd, const uint32_t* __restrict__ mask,
                          const int32_t* __restrict__ aoff, const int32_t* __restrict__ len,
                          int32_t n_seq, uint32_t* __restrict__ rc, uint32_t* __restrict__ mrc) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < n_seq; t += nwarps) {
        int64_t off = aoff[t];
        int32_t n = len[t];
        int32_t nw = (n + 15) >> 4;
        for (int32_t w = lane; w < nw; w += 32) {
            int32_t q = n - 16 * w - 16;
            uint32_t x = (uint32_t)fetch32(fwd, off + max(q, 0));
            int v = 16;
            if (q < 0) { v = 16 + q; x <<= 2 * (16 - v); }
            uint32_t r = ~rev2_32(x);
            if (v < 16) r &= (1u << (2 * v)) - 1u;
            rc[(off >> 4) + w] = r;
        }
        int32_t nm = (n + 31) >> 5;
        for (int32_t w = lane; w < nm; w += 32) {
            int32_t q = n - 32 * w - 32;
            uint32_t x = fetchmask32(mask, off + max(q, 0));
            int v = 32;
            if (q < 0) { v = 32 + q; x <<= (32 - v); }
            uint32_t r = __brev(x);
            if (v < 32) r |= ~((1u << v) - 1u);
            mrc[(off >> 5) + w] = r;
        }
    }
}

// Canonical-16-mer hash at arena position p (0xFFFFFFFF = contains a masked base).
// *pal = the 16-mer equals its own reverse complement.
__device__ __forceinline__ uint32_t kmer_hash(const uint32_t* fwd, const uint32_t* mask, int64_t p,
                                              uint32_t* orient, uint32_t* pal) {
    if (fetchmask32(mask, p) & 0xFFFFu) return 0xFFFFFFFFu;
    uint32_t f = (uint32_t)fetch32(fwd, p);
    uint32_t r = rev2_32(~f);
    uint32_t hf = rcq_mix32(f), hr = rcq_mix32(r);
    *orient = hr < hf;
    *pal = hr == hf;
    uint32_t h = min(hf, hr);
    return min(h, 0xFFFFFFFEu);
}

// reverse complement of 13 bases held in the low 26 bits
__device__ __forceinline__ uint32_t rc13(uint32_t x) { return rev2_32(~x & 0x03FFFFFFu) >> 6; }

// Flank context of the index entry at transcript position p (arena offset off, transcript length n,
// orientation o of its canonical 16-mer): the 13 bases before and after the 16-mer, read on the
// canonical strand.  Word layout: bits 0..25 bases, bit 30 (low word only) = palindromic 16-mer,
// bit 31 = flank not available (runs off the transcript or holds a masked base).
// Two entries of the same key whose "incoming" flanks are equal and available lie in an exact run
// that started >= 13 bases earlier, so they can never be the leftmost joint anchor of that run:
// k_seed_probe drops such key matches without touching the sequences (seed_kernels.cuh).
__device__ __forceinline__ uint64_t flank_ctx(const uint32_t* fwd, const uint32_t* mask, int64_t off, int32_t p,
                                              int32_t n, uint32_t o, uint32_t pal) {
    const uint32_t M26 = 0x03FFFFFFu, M13 = 0x1FFFu;
    uint32_t L = 0x80000000u, R = 0x80000000u;
    if (p >= RCQ_WIN && !(fetchmask32(mask, off + p - RCQ_WIN) & M13))
        L = (uint32_t)fetch32(fwd, off + p - RCQ_WIN) & M26;
    if (p + RCQ_KMER + RCQ_WIN <= n && !(fetchmask32(mask, off + p + RCQ_KMER) & M13))
        R = (uint32_t)fetch32(fwd, off + p + RCQ_KMER) & M26;
    uint32_t cl, cr;
    if (!o) { cl = L; cr = R; }
    else {
        cl = (R >> 31) ? R : rc13(R);
        cr = (L >> 31) ? L : rc13(L);
    }
    if (pal) cl |= 0x40000000u;
    return ((uint64_t)cr << 32) | cl;
}

// Window-minimizer selection (window = 13 consecutive 16-mers = one 28-mer).  Position p is
// selected iff its hash equals the minimum of SOME window containing it (all tied minima are
// kept, so the choice is symmetric under reverse complement).  Warp per transcript; selected
// entries are staged in a per-warp shared-memory buffer and appended with ONE atomic per flush
// (a single global counter serialises at ~1 atomic/ns, which bounded the per-step version).
// The 32 positions of a warp step are one aligned word of the `sel` bitmap: plain store of the ballot.
// Entries past `cap` are counted but not written (the host then re-runs with the exact size).
#define MZ_WARPS 8
#define MZ_BUF 256
__global__ void __launch_bounds__(32 * MZ_WARPS) k_minimizers(
    const uint32_t* __restrict__ fwd, const uint32_t* __restrict__ mask, const int32_t* __restrict__ aoff,
    const int32_t* __restrict__ len, int32_t n_seq, unsigned long long* __restrict__ counter,
    uint64_t* __restrict__ out, uint64_t* __restrict__ out_ctx, unsigned long long cap,
    uint32_t* __restrict__ selmap) {
    __shared__ uint32_t sh[MZ_WARPS][256 + 8];
    __shared__ uint8_t so[MZ_WARPS][256 + 8];
    __shared__ uint64_t be[MZ_WARPS][MZ_BUF], bc[MZ_WARPS][MZ_BUF];
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp = blockIdx.x * MZ_WARPS + wib;
    const int nwarps = gridDim.x * MZ_WARPS;
    const int T = 224, H = RCQ_WIN - 1;  // tile and halo
    int nbuf = 0;                        // warp-uniform fill of be/bc
    auto flush = [&]() {
        if (!nbuf) return;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(counter, (unsigned long long)nbuf);
        base = __shfl_sync(0xffffffffu, base, 0);
        __syncwarp();
        for (int i = lane; i < nbuf; i += 32)
            if (base + i < cap) { out[base + i] = be[wib][i]; out_ctx[base + i] = bc[wib][i]; }
        __syncwarp();
        nbuf = 0;
    };
    for (int t = warp; t < n_seq; t += nwarps) {
        const int64_t off = aoff[t];
        const int32_t n = len[t];
        const int32_t nk = n - RCQ_KMER + 1;
        if (nk < RCQ_WIN) continue;
        for (int32_t c = 0; c < nk; c += T) {
            __syncwarp();
            for (int i = lane; i < T + 2 * H; i += 32) {
                int32_t p = c - H + i;
                uint32_t o = 0, pal = 0, h = 0xFFFFFFFFu;
                if (p >= 0 && p < nk) h = kmer_hash(fwd, mask, off + p, &o, &pal);
                sh[wib][i] = h;
                so[wib][i] = (uint8_t)(o | (pal << 1));
            }
            __syncwarp();
            for (int i0 = 0; i0 < T; i0 += 32) {
                const int i = i0 + lane;
                const int32_t p = c + i;
                bool sel = false;
                uint32_t h = 0xFFFFFFFFu;
                if (p < nk) {
                    h = sh[wib][i + H];
                    if (h != 0xFFFFFFFFu) {
                        int a = 0, b = 0;
                        int amax = min(H, p), bmax = min(H, nk - 1 - p);
                        while (a < amax && sh[wib][i + H - a - 1] >= h) a++;
                        while (b < bmax && sh[wib][i + H + b + 1] >= h) b++;
                        sel = (H - b) <= a;  // some window [p-j, p-j+12], j<=a, 12-j<=b
                    }
                }
                const unsigned m = __ballot_sync(0xffffffffu, sel);
                if (c + i0 < nk && lane == 0) selmap[(off + c + i0) >> 5] = m;
                if (m) {
                    if (nbuf + 32 > MZ_BUF) flush();
                    if (sel) {
                        const uint32_t oo = so[wib][i + H];
                        const int slot = nbuf + __popc(m & ((1u << lane) - 1u));
                        be[wib][slot] = ((unsigned long long)rcq_index_key(h) << 32) |
                                        ((unsigned long long)(off + p) << 1) | (oo & 1u);
                        bc[wib][slot] = flank_ctx(fwd, mask, off, p, n, oo & 1u, oo >> 1);
                    }
                    nbuf += __popc(m);
                }
            }
        }
    }
    flush();
}

// blk2tx / ol tables: warp per transcript
__global__ void k_tx_tables(const int32_t* __restrict__ aoff, const int32_t* __restrict__ len, int32_t n_seq,
                            int32_t* __restrict__ blk2tx, int2* __restrict__ ol) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < n_seq; t += nwarps) {
        int32_t o = aoff[t], n = len[t];
        if (lane == 0) ol[t] = make_int2(o, n);
        for (int32_t b = lane; b < ((n + 31) >> 5); b += 32) blk2tx[(o >> 5) + b] = t;
    }
}

__global__ void k_build_dir(const uint64_t* __restrict__ index, int64_t n, uint32_t* __restrict__ dir) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nb = 1u << RCQ_DIRBITS;
    if (i > n) return;
    uint32_t lo = (i == 0) ? 0u : (uint32_t)(index[i - 1] >> (64 - RCQ_DIRBITS)) + 1u;
    uint32_t hi = (i == n) ? nb : (uint32_t)(index[i] >> (64 - RCQ_DIRBITS));
    for (uint32_t b = lo; b <= hi; b++) dir[b] = (uint32_t)i;
}

// index entries sorted by key (bits 32..63) with their flank contexts as payload
static int sort_index(rcq_ctx* ctx, uint64_t** keys, uint64_t** vals, int64_t n) {
    if (n <= 1) return 0;
    uint64_t *k2 = nullptr, *v2 = nullptr;
    RC(dalloc(ctx, &k2, n));
    RC(dalloc(ctx, &v2, n));
    cub::DoubleBuffer<uint64_t> dk(*keys, k2), dv(*vals, v2);
    size_t bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, dk, dv, (int)n, 32, 64, ctx->stream));
    void* ws = nullptr;
    CU(cudaMallocFromPoolAsync(&ws, bytes, ctx->pool, ctx->stream));
    CU(cub::DeviceRadixSort::SortPairs(ws, bytes, dk, dv, (int)n, 32, 64, ctx->stream));
    ctx->launches += 32 / 8 + 1;
    cudaFreeAsync(ws, ctx->stream);
    uint64_t *ko = dk.Current(), *vo = dv.Current();
    dfree(ctx, ko == *keys ? k2 : *keys);
    dfree(ctx, vo == *vals ? v2 : *vals);
    *keys = ko; *vals = vo;
    return 0;
}

// Pack + reverse complement + minimizer index for one sample (stream-ordered).
int rcq_sample_build(rcq_ctx* ctx, rcq_sample* s, const uint8_t* d_ascii, int64_t ascii_len,
                     const int64_t* d_starts) {
    int64_t words = s->arena / 16, mwords = s->arena / 32;
    RC(dalloc(ctx, &s->fwd, words));
    RC(dalloc(ctx, &s->rc, words));
    RC(dalloc(ctx, &s->mask, mwords));
    RC(dalloc(ctx, &s->mask_rc, mwords));
    RC(dalloc(ctx, &s->dir, (1ll << RCQ_DIRBITS) + 1));
    CU(cudaMemsetAsync(s->fwd, 0, words * 4, ctx->stream));
    CU(cudaMemsetAsync(s->rc, 0, words * 4, ctx->stream));
    CU(cudaMemsetAsync(s->mask, 0xFF, mwords * 4, ctx->stream));
    CU(cudaMemsetAsync(s->mask_rc, 0xFF, mwords * 4, ctx->stream));
    RC(dalloc(ctx, &s->sel, mwords));
    RC(dalloc(ctx, &s->blk2tx, mwords));
    RC(dalloc(ctx, &s->ol, s->n_seq));
    CU(cudaMemsetAsync(s->blk2tx, 0, mwords * 4, ctx->stream));
    CU(cudaMemsetAsync(s->sel, 0, mwords * 4, ctx->stream));
    unsigned long long* ctr = (unsigned long long*)ctx->d_ctr;
    CU(cudaMemsetAsync(ctr, 0, 2 * sizeof(unsigned long long), ctx->stream));
    int blocks = RCQ_SMS * 8;
    if (s->n_seq) {
        LAUNCH(ctx, k_pack, blocks, 256, 0, d_ascii, ascii_len, d_starts, s->off, s->len, s->n_seq, s->fwd,
               s->mask, ctr);
        LAUNCH(ctx, k_tx_tables, blocks, 256, 0, s->off, s->len, s->n_seq, s->blk2tx, s->ol);
        LAUNCH(ctx, k_revcomp, blocks, 256, 0, s->fwd, s->mask, s->off, s->len, s->n_seq, s->rc,
               s->mask_rc);
    }
    // minimizer entries: one pass into a buffer sized for the usual density (2/14 per base);
    // a denser sample is re-run once with the exact count
    int64_t cap = s->total / 5 + 4096;
    for (int attempt = 0; attempt < 2; attempt++) {
        RC(dalloc(ctx, &s->index, cap));
        RC(dalloc(ctx, &s->fctx, cap));
        CU(cudaMemsetAsync(ctr + 1, 0, sizeof(unsigned long long), ctx->stream));
        if (s->n_seq)
            LAUNCH(ctx, k_minimizers, blocks, 32 * MZ_WARPS, 0, s->fwd, s->mask, s->off, s->len, s->n_seq, ctr + 1,
                   s->index, s->fctx, (unsigned long long)cap, s->sel);
        CU(cudaMemcpyAsync(ctx->h_pin, ctr, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        s->n_masked = ctx->h_pin[0];
        s->n_index = ctx->h_pin[1];
        if (s->n_index <= cap) break;
        dfree(ctx, s->index); s->index = nullptr;
        dfree(ctx, s->fctx); s->fctx = nullptr;
        cap = s->n_index;
    }
    // sorted by key only (bits 32..63): entries of one key stay in append order, which no result
    // depends on (seeds are re-sorted by transcript and diagonal)
    if (s->n_index) RC(sort_index(ctx, &s->index, &s->fctx, s->n_index));
    LAUNCH(ctx, k_build_dir, grid_for(s->n_index + 1, 256), 256, 0, s->index, s->n_index, s->dir);
    return 0;
}
