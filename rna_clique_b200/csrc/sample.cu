// sample.cu — kernels (1) and (2): 2-bit packing of a sample's top-gene transcripts and the
// sorted minimizer seed index.  Replaces `makeblastdb` per sample
// (reference find_all_pairs.py:120-159).  HBM-bound streaming kernels; algorithmic bytes:
//   pack   : 1 B/nt read + 0.25 B/nt written (+0.25 rc +2x1/32 masks)
//   index  : 0.25 B/nt read, 8 B per selected minimizer written, then an 8-B-key radix sort.
#include <cub/cub.cuh>
#include <mutex>
#include "sample.cuh"
#include "dust_kernels.cuh"

#define GUARD_FRONT 64
#define GUARD_BACK 128

__device__ __forceinline__ uint32_t base_code(uint32_t c) {  // 0..3, or 4 = not ACGT(U)
    uint32_t u = c & 0xDFu;
    return u == 'A' ? 0u : u == 'C' ? 1u : u == 'G' ? 2u : (u == 'T' || u == 'U') ? 3u : 4u;
}

// Four ASCII letters (one per byte of w) -> bits 0..7: their 2-bit codes, bits 8..11: "not A, C, G, T or U"
// (either case).  With u = letter & 0xDF the code is Gray-decoded from bits 1-2 of u (A 00, C 01, G 11,
// T/U 10), and the letter is valid iff the remaining bits say 0x41 with bits 1-2 != 10 (A, C, G) or
// 0x50 / 0x51 with bits 1-2 == 10 (T, U).  Byte-parallel: no carries cross byte lanes.
__device__ __forceinline__ uint32_t pack4(uint32_t w) {
    const uint32_t u = w & 0xDFDFDFDFu;
    const uint32_t r = (u >> 1) & 0x03030303u;
    const uint32_t code = r ^ ((r >> 1) & 0x01010101u);
    const uint32_t h = u & 0xF9F9F9F9u;
    // per-byte zero tests (exact: 0x80 in a byte iff that byte of x is zero)
#define ZB(x) (~((((x) & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | (x) | 0x7F7F7F7Fu))
    const uint32_t is41 = ZB(h ^ 0x41414141u);
    const uint32_t is5x = ZB((h & 0xFEFEFEFEu) ^ 0x50505050u);       // 0x50 or 0x51
    const uint32_t r2 = ZB(r ^ 0x02020202u);
#undef ZB
    const uint32_t valid = (is41 & ~r2) | (is5x & r2);              // 0x80 per valid byte
    const uint32_t bad = (~valid & 0x80808080u) >> 7;               // 0x01 per invalid byte
    // gather: byte k of code -> bits 2k..2k+1, byte k of bad -> bit k (multiply = shifted adds, no overlaps)
    const uint32_t c8 = (code * 0x01041040u) >> 24;
    const uint32_t b4 = (bad * 0x10204080u) >> 28;
    return (c8 & 0xFFu) | ((b4 & 0xFu) << 8);
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
                const int lim = min(16, n - b0);
                // four letters per 32-bit lane of arithmetic (pack4): 8 code bits + 4 "not ACGTU" bits
                const uint32_t q[4] = {(uint32_t)A, (uint32_t)(A >> 32), (uint32_t)B, (uint32_t)(B >> 32)};
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t r = pack4(q[k]);
                    word |= (r & 0xFFu) << (8 * k);
                    mbits |= (r >> 8) << (4 * k);
                }
                const uint32_t inside = lim >= 16 ? 0xFFFFu : ((1u << lim) - 1u);
                masked += __popc(mbits & inside);
                mbits |= ~inside & 0xFFFFu;
                // codes under a mask bit are zero, as before
                uint32_t keep2 = ~mbits & 0xFFFFu;                      // spread 16 bits to 16 bit pairs
                keep2 = (keep2 | (keep2 << 8)) & 0x00FF00FFu;
                keep2 = (keep2 | (keep2 << 4)) & 0x0F0F0F0Fu;
                keep2 = (keep2 | (keep2 << 2)) & 0x33333333u;
                keep2 = (keep2 | (keep2 << 1)) & 0x55555555u;
                word &= keep2 | (keep2 << 1);
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
// kept, so the choice is symmetric under reverse complement); windows lie inside the transcript.
//
// Warp per transcript, tiles of 480 positions: lane l hashes the 16 word-aligned positions
// [c - 16 + 16 l, +16) from four 2-bit words held in registers (lanes 0 and 31 are the halo of the
// tile), parks the hashes in shared memory, reads back its 12 + 16 + 12 neighbourhood and decides
// its 16 positions with two branch-free doubling passes in registers:
//   w[t] = min of the 13 hashes of window t (0 for a window that leaves the transcript),
//   sel  = max of the 13 windows containing p == hash(p).
// Entries and their flank contexts (sample.cu: flank_ctx semantics) are staged in a per-warp
// buffer and appended with one atomic per flush; two lanes' decisions form one aligned word of `sel`.
// Entries past `cap` are counted but not written (the host then re-runs with the exact size).
#define MZ_WARPS 4
#define MZ_BUF 576
#define MZ_TILE 480
#define MZ_STRIDE 20   // padded words per lane in the hash exchange buffer (bank-conflict-free LDS.128)
__global__ void __launch_bounds__(32 * MZ_WARPS) k_minimizers(
    const uint32_t* __restrict__ fwd, const uint32_t* __restrict__ mask, const int32_t* __restrict__ aoff,
    const int32_t* __restrict__ len, int32_t n_seq, unsigned long long* __restrict__ counter,
    uint64_t* __restrict__ out, uint64_t* __restrict__ out_ctx, unsigned long long cap,
    uint32_t* __restrict__ selmap) {
    __shared__ __align__(16) uint32_t hx[MZ_WARPS][32 * MZ_STRIDE];
    __shared__ uint64_t be[MZ_WARPS][MZ_BUF], bc[MZ_WARPS][MZ_BUF];
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp = blockIdx.x * MZ_WARPS + wib;
    const int nwarps = gridDim.x * MZ_WARPS;
    const uint32_t M26 = 0x03FFFFFFu;
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
        for (int32_t c = 0; c < nk; c += MZ_TILE) {
            const int32_t q0 = c - 16 + 16 * lane;          // first position of this lane (may be < 0 or >= nk)
            // ---- bases [q0-16, q0+48) as four words, mask bits of the same span
            uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            uint64_t M64 = ~0ull;
            const bool live = q0 + 16 > 0 && q0 < nk;       // some position of the lane is a k-mer start
            if (live) {
                const int64_t w = (off + q0 - 16) >> 4;     // >= 2: the arena has a 64-base front guard
                a0 = __ldg(fwd + w); a1 = __ldg(fwd + w + 1); a2 = __ldg(fwd + w + 2); a3 = __ldg(fwd + w + 3);
                const int64_t mw = (off + q0 - 16) >> 5;
                const uint32_t m0 = __ldg(mask + mw), m1 = __ldg(mask + mw + 1), m2 = __ldg(mask + mw + 2);
                if ((off + q0 - 16) & 16) M64 = ((uint64_t)(m0 >> 16)) | ((uint64_t)m1 << 16) | ((uint64_t)m2 << 48);
                else M64 = (uint64_t)m0 | ((uint64_t)m1 << 32);
            }
            // ---- 16 canonical hashes
            uint32_t h[16];
            uint32_t obits = 0, pbits = 0;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const int32_t p = q0 + j;
                const uint32_t f = __funnelshift_r(a1, a2, 2 * j);
                const uint32_t r = rev2_32(~f);
                const uint32_t hf = rcq_mix32(f), hr = rcq_mix32(r);
                uint32_t v = min(min(hf, hr), 0xFFFFFFFEu);
                if (p < 0 || p >= nk || ((uint32_t)(M64 >> (16 + j)) & 0xFFFFu)) v = 0xFFFFFFFFu;
                h[j] = v;
                obits |= (uint32_t)(hr < hf) << j;
                pbits |= (uint32_t)(hr == hf) << j;
            }
            __syncwarp();
            uint4* mine = (uint4*)(hx[wib] + lane * MZ_STRIDE);
            mine[0] = make_uint4(h[0], h[1], h[2], h[3]);
            mine[1] = make_uint4(h[4], h[5], h[6], h[7]);
            mine[2] = make_uint4(h[8], h[9], h[10], h[11]);
            mine[3] = make_uint4(h[12], h[13], h[14], h[15]);
            __syncwarp();
            // ---- neighbourhood v[0..40) = hashes of positions q0-12 .. q0+27
            uint32_t v[40];
            {
                const uint4* L = (const uint4*)(hx[wib] + ((lane + 31) & 31) * MZ_STRIDE);
                const uint4* R = (const uint4*)(hx[wib] + ((lane + 1) & 31) * MZ_STRIDE);
                const uint4 l1 = L[1], l2 = L[2], l3 = L[3], r0 = R[0], r1 = R[1], r2 = R[2];
                v[0] = l1.x; v[1] = l1.y; v[2] = l1.z; v[3] = l1.w; v[4] = l2.x; v[5] = l2.y; v[6] = l2.z; v[7] = l2.w;
                v[8] = l3.x; v[9] = l3.y; v[10] = l3.z; v[11] = l3.w;
#pragma unroll
                for (int j = 0; j < 16; j++) v[12 + j] = h[j];
                v[28] = r0.x; v[29] = r0.y; v[30] = r0.z; v[31] = r0.w; v[32] = r1.x; v[33] = r1.y; v[34] = r1.z; v[35] = r1.w;
                v[36] = r2.x; v[37] = r2.y; v[38] = r2.z; v[39] = r2.w;
            }
            // ---- w[t] = min(v[t..t+12]), t < 28, by doubling (in place)
#pragma unroll
            for (int i = 0; i < 39; i++) v[i] = min(v[i], v[i + 1]);       // spans of 2
#pragma unroll
            for (int i = 0; i < 37; i++) v[i] = min(v[i], v[i + 2]);       // 4
#pragma unroll
            for (int i = 0; i < 33; i++) v[i] = min(v[i], v[i + 4]);       // 8
#pragma unroll
            for (int i = 0; i < 28; i++) {
                uint32_t x = min(v[i], v[i + 5]);                          // 13
                const int32_t ps = q0 - 12 + i;                            // window start
                v[i] = (ps < 0 || ps > nk - RCQ_WIN) ? 0u : x;
            }
            // ---- x[j] = max(w[j..j+12]), j < 16
#pragma unroll
            for (int i = 0; i < 27; i++) v[i] = max(v[i], v[i + 1]);
#pragma unroll
            for (int i = 0; i < 25; i++) v[i] = max(v[i], v[i + 2]);
#pragma unroll
            for (int i = 0; i < 21; i++) v[i] = max(v[i], v[i + 4]);
            uint32_t sel = 0;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const uint32_t x = max(v[j], v[j + 5]);
                sel |= (uint32_t)(h[j] != 0xFFFFFFFFu && x == h[j]) << j;
            }
            if (lane == 0 || lane == 31) sel = 0;          // halo lanes decide nothing
            // ---- sel bitmap: lanes (odd, odd+1) make one aligned 32-position word
            {
                const uint32_t hi = __shfl_down_sync(0xffffffffu, sel, 1);
                if ((lane & 1) && lane < 31 && q0 < nk) selmap[(off + q0) >> 5] = sel | (hi << 16);
            }
            // ---- append
            const int cnt = __popc(sel);
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int x = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += x;
            }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            if (total) {
                if (nbuf + total > MZ_BUF) flush();
                int slot = nbuf + incl - cnt;
                // one iteration per selected position of the lane (~2 of 16): dynamic j, so the hash is
                // re-read from the exchange buffer and the flank words are picked by predicate
                const uint32_t* myh = hx[wib] + lane * MZ_STRIDE;
                for (uint32_t rest = sel; rest; rest &= rest - 1) {
                    const int j = __ffs(rest) - 1;
                    const int32_t p = q0 + j;
                    const uint32_t o = (obits >> j) & 1u;
                    // flanks: 13 bases before (window offset 3+j) and after (window offset 32+j) the 16-mer
                    uint32_t Lf = 0x80000000u, Rf = 0x80000000u;
                    if (p >= RCQ_WIN && !((uint32_t)(M64 >> (3 + j)) & 0x1FFFu)) {
                        const int bo = 3 + j;
                        const uint32_t lo = bo < 16 ? a0 : a1, hi = bo < 16 ? a1 : a2;
                        Lf = __funnelshift_r(lo, hi, 2 * (bo & 15)) & M26;
                    }
                    if (p + RCQ_KMER + RCQ_WIN <= n && !((uint32_t)(M64 >> (32 + j)) & 0x1FFFu))
                        Rf = __funnelshift_r(a2, a3, 2 * j) & M26;
                    uint32_t cl, cr;
                    if (!o) { cl = Lf; cr = Rf; }
                    else {
                        cl = (Rf >> 31) ? Rf : rc13(Rf);
                        cr = (Lf >> 31) ? Lf : rc13(Lf);
                    }
                    if ((pbits >> j) & 1u) cl |= 0x40000000u;
                    be[wib][slot] = ((unsigned long long)rcq_index_key(myh[j]) << 32) |
                                    ((unsigned long long)(off + p) << 1) | o;
                    bc[wib][slot] = ((uint64_t)cr << 32) | cl;
                    slot++;
                }
                nbuf += total;
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
static int sort_index(rcq_ctx* ctx, uint64_t* keys, uint64_t* vals, int64_t n) {
    if (n <= 1) return 0;
    uint64_t *k2 = nullptr, *v2 = nullptr;
    RC(talloc(ctx, &k2, n));
    RC(talloc(ctx, &v2, n));
    cub::DoubleBuffer<uint64_t> dk(keys, k2), dv(vals, v2);
    size_t bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, dk, dv, (int)n, 32, 64, ctx->stream));
    char* ws = nullptr;
    RC(talloc(ctx, &ws, (int64_t)bytes));
    CU(cub::DeviceRadixSort::SortPairs(ws, bytes, dk, dv, (int)n, 32, 64, ctx->stream));
    ctx->launches += 32 / 8 + 1;
    // four 8-bit passes end in the buffers they started from; copy back if a CUB version ever differs
    if (dk.Current() != keys) CU(cudaMemcpyAsync(keys, dk.Current(), n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    if (dv.Current() != vals) CU(cudaMemcpyAsync(vals, dv.Current(), n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
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
    RC(dalloc(ctx, &s->smask, mwords));
    RC(dalloc(ctx, &s->smask_rc, mwords));
    CU(cudaMemsetAsync(s->smask, 0, mwords * 4, ctx->stream));
    CU(cudaMemsetAsync(s->smask_rc, 0xFF, mwords * 4, ctx->stream));
    unsigned long long* ctr = (unsigned long long*)ctx->d_ctr;
    CU(cudaMemsetAsync(ctr, 0, 8 * sizeof(unsigned long long), ctx->stream));
    int blocks = RCQ_SMS * 8;
    if (s->n_seq) {
        LAUNCH(ctx, k_pack, blocks, 256, 0, d_ascii, ascii_len, d_starts, s->off, s->len, s->n_seq, s->fwd,
               s->mask, ctr);
        LAUNCH(ctx, k_tx_tables, blocks, 256, 0, s->off, s->len, s->n_seq, s->blk2tx, s->ol);
        LAUNCH(ctx, k_revcomp, blocks, 256, 0, s->fwd, s->mask, s->off, s->len, s->n_seq, s->rc,
               s->mask_rc);
        // seeding mask: N mask | symmetric DUST (what blastn's default query filter hides from seeding)
        if (ctx->dust_on) {
            uint8_t* trip = nullptr; uint32_t* flag = nullptr;
            RC(talloc(ctx, &trip, s->arena));
            RC(talloc(ctx, &flag, mwords));
            LAUNCH(ctx, k_dust_triplets, grid_for(words, 256), 256, 0, s->fwd, s->mask, s->arena, trip);
            LAUNCH(ctx, k_dust_flag, grid_for((s->arena + DUST_SEG - 1) / DUST_SEG, DUST_FLAG_THREADS), DUST_FLAG_THREADS, 0,
                   trip, s->arena, flag);
            LAUNCH(ctx, k_dust_eval, RCQ_SMS * 8, DUST_EVAL_THREADS, 0, trip, flag, s->arena, s->smask, ctr);
            tfree(ctx, trip); tfree(ctx, flag);
        }
        LAUNCH(ctx, k_dust_or, grid_for(mwords, 256), 256, 0, s->mask, s->smask, mwords, ctr);
        LAUNCH(ctx, k_revmask, blocks, 256, 0, s->smask, s->off, s->len, s->n_seq, s->smask_rc);
    }
    // minimizer entries: one pass into a buffer sized for the usual density (2/14 per base);
    // a denser sample is re-run once with the exact count
    int64_t cap = s->total / 5 + 4096;
    for (int attempt = 0; attempt < 2; attempt++) {
        RC(dalloc(ctx, &s->index, cap));
        RC(dalloc(ctx, &s->fctx, cap));
        CU(cudaMemsetAsync(ctr + 1, 0, sizeof(unsigned long long), ctx->stream));
        if (s->n_seq)
            LAUNCH(ctx, k_minimizers, blocks, 32 * MZ_WARPS, 0, s->fwd, s->smask, s->off, s->len, s->n_seq, ctr + 1,
                   s->index, s->fctx, (unsigned long long)cap, s->sel);
        RC(publish_counters(ctx, (const int64_t*)ctr, 5));
        s->n_masked = ctx->h_pin[0];
        s->n_index = ctx->h_pin[1];
        s->n_dust = ctx->h_pin[4];
        if (s->n_index <= cap) break;
        dfree(ctx, s->index); s->index = nullptr;
        dfree(ctx, s->fctx); s->fctx = nullptr;
        cap = s->n_index;
    }
    // sorted by key only (bits 32..63): entries of one key stay in append order, which no result
    // depends on (seeds are re-sorted by transcript and diagonal)
    if (s->n_index) RC(sort_index(ctx, s->index, s->fctx, s->n_index));
    LAUNCH(ctx, k_build_dir, grid_for(s->n_index + 1, 256), 256, 0, s->index, s->n_index, s->dir);
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// rcq_dust_mask_static — the DUST seeding-mask kernels of the sample build (k_dust_triplets, k_dust_flag,
// k_dust_eval, k_dust_or, launched exactly as above) for callers that must not allocate: their scratch (triplet
// bytes, flag words, counters) lives in the file-scope device array below instead of the context's pool.  No
// counterpart in the reference (include/rcq.h).  The array is held by
// every GPU that loads the library until the process exits; every call on a GPU shares it, so calls that overlap
// on one GPU race.
#define RCQ_DUST_STATIC_CTR_BYTES 256
#define RCQ_DUST_STATIC_FLAG_BYTES (RCQ_DUST_STATIC_MAX_BASES / 8)
#define RCQ_DUST_STATIC_BYTES (RCQ_DUST_STATIC_CTR_BYTES + RCQ_DUST_STATIC_FLAG_BYTES + RCQ_DUST_STATIC_MAX_BASES)
static __device__ __align__(256) uint8_t g_dust_static[RCQ_DUST_STATIC_BYTES];

#define RCQ_DUST_STATIC_MAX_DEVICES 64
static uint8_t* g_dust_static_ptr[RCQ_DUST_STATIC_MAX_DEVICES];   // device address of g_dust_static on each GPU
static std::mutex g_dust_static_mu;

namespace {
struct DeviceGuard {        // makes `device` current for the call and puts the caller's device back
    int prev = -1;
    bool changed = false;
    cudaError_t enter(int device) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev != device) { e = cudaSetDevice(device); changed = e == cudaSuccess; }
        return e;
    }
    ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
};
}  // namespace

// In host code `g_dust_static` is only a stand-in for the symbol: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static int dust_static_base(int device, uint8_t** base) {
    std::lock_guard<std::mutex> lock(g_dust_static_mu);
    if (!g_dust_static_ptr[device]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_dust_static));
        if (!p) return rcq_fail(RCQ_ERR_CUDA, "rcq_dust_mask_static: cudaGetSymbolAddress returned a null address");
        g_dust_static_ptr[device] = (uint8_t*)p;
    }
    *base = g_dust_static_ptr[device];
    return 0;
}

extern "C" int rcq_dust_mask_static(int device, const uint32_t* dev_fwd, const uint32_t* dev_mask, int64_t arena_bases,
                                    uint32_t* dev_smask, void* stream, int64_t* n_dust) {
    if (device < 0 || device >= RCQ_DUST_STATIC_MAX_DEVICES || !dev_fwd || !dev_mask || !dev_smask)
        return rcq_fail(RCQ_ERR_ARG, "rcq_dust_mask_static: bad argument");
    if (arena_bases <= 0 || arena_bases % 64)
        return rcq_fail(RCQ_ERR_ARG, "rcq_dust_mask_static: arena_bases must be a positive multiple of 64");
    if (arena_bases > RCQ_DUST_STATIC_MAX_BASES) {
        char a[32], b[32];
        snprintf(a, sizeof a, "%lld", (long long)arena_bases);
        snprintf(b, sizeof b, "%lld", (long long)RCQ_DUST_STATIC_MAX_BASES);
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_dust_mask_static: %s bases exceed the static scratch (at most %s bases per call)", a, b);
    }
    DeviceGuard guard;
    CU(guard.enter(device));
    uint8_t* base = nullptr;
    RC(dust_static_base(device, &base));
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long* ctr = (unsigned long long*)base;
    uint32_t* flag = (uint32_t*)(base + RCQ_DUST_STATIC_CTR_BYTES);
    uint8_t* trip = base + RCQ_DUST_STATIC_CTR_BYTES + RCQ_DUST_STATIC_FLAG_BYTES;
    const int64_t words = arena_bases / 16, mwords = arena_bases / 32;
    // the array keeps the last call's contents: counters and the output (k_dust_eval ORs into it) start at zero;
    // triplets and flags are written in full for [0, arena_bases) before they are read
    CU(cudaMemsetAsync(ctr, 0, RCQ_DUST_STATIC_CTR_BYTES, st));
    CU(cudaMemsetAsync(flag, 0, (size_t)mwords * 4, st));
    CU(cudaMemsetAsync(dev_smask, 0, (size_t)mwords * 4, st));
    k_dust_triplets<<<grid_for(words, 256), 256, 0, st>>>(dev_fwd, dev_mask, arena_bases, trip);
    CU(cudaGetLastError());
    k_dust_flag<<<grid_for((arena_bases + DUST_SEG - 1) / DUST_SEG, DUST_FLAG_THREADS), DUST_FLAG_THREADS, 0, st>>>(trip, arena_bases, flag);
    CU(cudaGetLastError());
    k_dust_eval<<<RCQ_SMS * 8, DUST_EVAL_THREADS, 0, st>>>(trip, flag, arena_bases, dev_smask, ctr);
    CU(cudaGetLastError());
    k_dust_or<<<grid_for(mwords, 256), 256, 0, st>>>(dev_mask, dev_smask, mwords, ctr);
    CU(cudaGetLastError());
    if (n_dust) {
        unsigned long long v = 0;
        CU(cudaMemcpyAsync(&v, ctr + 4, sizeof v, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        *n_dust = (int64_t)v;
    }
    return 0;
}
