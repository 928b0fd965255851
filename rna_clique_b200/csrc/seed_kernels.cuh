// seed_kernels.cuh — kernel (3a): minimizer join -> exact 28-mer runs -> X-drop ungapped HSPs.
//
// For one unordered sample pair (A = sample 1 on its plus strand, B = sample 2 on both strands)
// the kernels find every maximal exact-match run of >= 28 bases between a transcript of A and a
// transcript of B (the megablast seed condition, blastn word_size 28) and extend each run without
// gaps under an X-drop of 10 (20 bits).  Any 28-mer window contains the window minimum of its 13
// canonical 16-mers, and both samples index every position that is a window minimum, so joining
// the two sorted minimizer lists on the key finds at least one anchor inside every such run.
// Exactly one anchor per run survives (the leftmost position indexed on both sides).
//
// Roofline: HBM/L2-bound.  Algorithmic bytes per pair: 8 B per index entry of A streamed +
// 4 B directory + 8 B entry of B per probe (sorted probes => sequential), then 2-bit sequence
// words for verification (L2-resident: 3 arenas of L/4 bytes).
#pragma once
#include "sample.cuh"
#include "search.cuh"

// transcript containing arena position p: last t with off[t] <= p
__device__ __forceinline__ int32_t find_tx(const int32_t* __restrict__ off, int32_t n, int32_t p) {
    int32_t lo = 0, hi = n;  // invariant: off[lo] <= p < off[hi]
    while (hi - lo > 1) {
        int32_t mid = (lo + hi) >> 1;
        if (__ldg(off + mid) <= p) lo = mid; else hi = mid;
    }
    return lo;
}

// number of consecutive matching base pairs going forward from (pa, pb), at most lim
__device__ __forceinline__ int32_t run_fwd(const SeqArena& A, int64_t pa, const SeqArena& B, int64_t pb,
                                           int32_t lim) {
    int32_t n = 0;
    while (n < lim) {
        uint32_t m = match32(A, pa + n, B, pb + n);
        int32_t z = (m == 0xFFFFFFFFu) ? 32 : (__ffs(~m) - 1);
        if (z > lim - n) z = lim - n;
        n += z;
        if (z < 32) break;
    }
    return n;
}
// same going backward from (pa, pb) inclusive
__device__ __forceinline__ int32_t run_back(const SeqArena& A, int64_t pa, const SeqArena& B, int64_t pb,
                                            int32_t lim) {
    int32_t n = 0;
    while (n < lim) {
        uint32_t m = match32_back(A, pa - n, B, pb - n);
        int32_t z = (m == 0xFFFFFFFFu) ? 32 : (__ffs(~m) - 1);
        if (z > lim - n) z = lim - n;
        n += z;
        if (z < 32) break;
    }
    return n;
}

// ---- phase 1: probe.  Thread per index entry of A; both indexes are sorted by the same key, so
// neighbouring threads read neighbouring directory buckets and entries of B (coalesced, L2
// friendly).  A key match whose incoming 13-base flanks are equal in both samples lies >= 13 bases
// inside a shared exact run, which always holds an earlier joint anchor (see k_seed_verify): it
// is dropped here from the 8-byte flank contexts alone (≈ 90 % of all key matches between
// orthologs), so only run starts and chance matches reach the sequence-touching verify kernel.
// Survivors become candidates (entry of A, entry of B), appended with one atomic per BLOCK.
// counters[1] = #candidates (may exceed cap: the host re-runs with a larger list),
// counters[2] = #key matches examined.
__device__ __forceinline__ bool probe_keep(uint64_t cb, uint32_t oA, uint32_t want, bool usable) {
    const uint32_t other = oA ? (uint32_t)(cb >> 32) : ((uint32_t)cb & 0xBFFFFFFFu);
    return !(usable && other == want);
}
// Block-tiled merge join.  A block owns PROBE_TA consecutive entries of A (registers, 4 per thread);
// their key span maps through B's directory to one contiguous range of B, which is streamed through
// shared memory PROBE_TB entries (+ contexts) at a time with coalesced 8-byte loads.  Every global
// access of the kernel is therefore a sequential stream (16 B per entry of A and of B, 8 B per
// surviving candidate); the per-entry lookup is a binary search in shared memory.
#ifndef PROBE_THREADS
#define PROBE_THREADS 64
#endif
#define PROBE_TA (4 * PROBE_THREADS)
#define PROBE_TB (6 * PROBE_THREADS)   // > PROBE_TA: the B range of a tile is ~PROBE_TA entries plus two buckets of slack
__global__ void __launch_bounds__(PROBE_THREADS) k_seed_probe(
    const uint64_t* __restrict__ idxA, const uint64_t* __restrict__ ctxA, int64_t nIdxA,
    const uint64_t* __restrict__ idxB, const uint64_t* __restrict__ ctxB, const uint32_t* __restrict__ dirB,
    uint2* __restrict__ cand, int64_t cap, unsigned long long* __restrict__ counters) {
    __shared__ uint64_t sB[PROBE_TB], sC[PROBE_TB];
    __shared__ uint32_t wsum[PROBE_THREADS / 32], wseen[PROBE_THREADS / 32];
    __shared__ unsigned long long bbase;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t a0 = (int64_t)blockIdx.x * PROBE_TA;
    const int64_t a1 = min(a0 + (int64_t)PROBE_TA, nIdxA);
    // thread t owns the 4 CONSECUTIVE entries a0 + 4t .. +3 (ascending keys): one binary search for the
    // first, the others continue the scan from where the previous one stopped
    uint32_t key[4], lowA[4], want[4];
    bool usable[4], have[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int64_t e = a0 + 4 * threadIdx.x + r;
        have[r] = e < a1;
        key[r] = 0xFFFFFFFFu; lowA[r] = 0; want[r] = 0; usable[r] = false;
        if (have[r]) {
            const uint64_t ka = __ldg(idxA + e), ca = __ldg(ctxA + e);
            key[r] = (uint32_t)(ka >> 32);
            lowA[r] = (uint32_t)ka;
            want[r] = (lowA[r] & 1u) ? (uint32_t)(ca >> 32) : ((uint32_t)ca & 0xBFFFFFFFu);
            usable[r] = !(want[r] >> 31) && !((uint32_t)ca & 0x40000000u);
        }
    }
    const uint32_t kfirst = (uint32_t)(__ldg(idxA + a0) >> 32), klast = (uint32_t)(__ldg(idxA + a1 - 1) >> 32);
    const uint32_t blo = __ldg(dirB + (kfirst >> (32 - RCQ_DIRBITS))), bhi = __ldg(dirB + (klast >> (32 - RCQ_DIRBITS)) + 1);
    for (uint32_t b0 = blo; b0 < bhi; b0 += PROBE_TB) {
        const int nb = (int)min((uint32_t)PROBE_TB, bhi - b0);
        for (int i = threadIdx.x; i < nb; i += PROBE_THREADS) { sB[i] = __ldg(idxB + b0 + i); sC[i] = __ldg(ctxB + b0 + i); }
        __syncthreads();
        // ---- count survivors of this chunk
        int pos[4];
        uint32_t cnt = 0, seen = 0;
        {
            int lo = 0, hi = have[0] ? nb : 0;        // first entry of the chunk with key >= key[0]
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if ((uint32_t)(sB[mid] >> 32) < key[0]) lo = mid + 1; else hi = mid;
            }
            int k = have[0] ? lo : nb;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                if (!have[r]) { pos[r] = nb; continue; }
                while (k < nb && (uint32_t)(sB[k] >> 32) < key[r]) k++;
                pos[r] = k;
                // equal keys of A share the run of B: do not advance k past it
                for (int q = k; q < nb && (uint32_t)(sB[q] >> 32) == key[r]; q++) {
                    seen++;
                    cnt += probe_keep(sC[q], lowA[r] & 1u, want[r], usable[r]);
                }
            }
        }
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        for (int o = 16; o; o >>= 1) seen += __shfl_down_sync(0xffffffffu, seen, o);
        if (lane == 31) wsum[wib] = incl;
        if (lane == 0) wseen[wib] = seen;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t tot = 0, sn = 0;
            for (int w = 0; w < PROBE_THREADS / 32; w++) { uint32_t v = wsum[w]; wsum[w] = tot; tot += v; sn += wseen[w]; }
            bbase = tot ? atomicAdd(counters + 1, (unsigned long long)tot) : 0ull;
            if (sn) atomicAdd(counters + 2, (unsigned long long)sn);
        }
        __syncthreads();
        if (cnt) {
            unsigned long long slot = bbase + wsum[wib] + (incl - cnt);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                for (int k = pos[r]; k < nb && (uint32_t)(sB[k] >> 32) == key[r]; k++) {
                    if (!probe_keep(sC[k], lowA[r] & 1u, want[r], usable[r])) continue;
                    if ((int64_t)slot < cap) cand[slot] = make_uint2(lowA[r], (uint32_t)sB[k]);
                    slot++;
                }
            }
        }
        __syncthreads();
    }
}

// ---- between probe and verify: candidates leave the probe in index-key order, i.e. at random positions of both
// samples, and the verify gathers about ten 32-byte sectors per candidate.  Three small kernels put them into
// CAND_BUCKETS buckets by their position in sample A (counting sort on the top bits of pA; order inside a bucket is
// whatever the atomics give, no result depends on it), so that neighbouring threads of the verify read neighbouring
// bases, masks and index bits of A - and of B wherever the two samples list their genes in the same order.  The
// candidate count stays on the device (counters[1]).
#define CAND_BUCKET_BITS 12
#define CAND_BUCKETS (1 << CAND_BUCKET_BITS)
#define CAND_PAD 32            // one counter per 128-byte line: atomics of different buckets never share a line
__global__ void __launch_bounds__(256) k_cand_hist(const uint2* __restrict__ cand, int64_t cand_cap,
                                                   const unsigned long long* __restrict__ counters, int shift,
                                                   uint32_t* __restrict__ hist) {
    __shared__ uint32_t h[CAND_BUCKETS];
    for (int i = threadIdx.x; i < CAND_BUCKETS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const int64_t n_cand = min((int64_t)counters[1], cand_cap);
    for (int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_cand; x += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&h[(__ldg(&cand[x].x) >> 1) >> shift], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < CAND_BUCKETS; i += blockDim.x)
        if (h[i]) atomicAdd(hist + (size_t)i * CAND_PAD, h[i]);
}
// counts -> exclusive start offsets, one block of 1024 threads
__global__ void __launch_bounds__(1024) k_cand_scan(uint32_t* __restrict__ hist) {
    __shared__ uint32_t wsum[32];
    constexpr int PER = CAND_BUCKETS / 1024;
    uint32_t v[PER], tot = 0;
#pragma unroll
    for (int r = 0; r < PER; r++) { v[r] = hist[(size_t)(threadIdx.x * PER + r) * CAND_PAD]; tot += v[r]; }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (w == 0) {
        uint32_t x = wsum[lane], i2 = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, i2, o); if (lane >= o) i2 += y; }
        wsum[lane] = i2 - x;
    }
    __syncthreads();
    uint32_t run = wsum[w] + incl - tot;
#pragma unroll
    for (int r = 0; r < PER; r++) { hist[(size_t)(threadIdx.x * PER + r) * CAND_PAD] = run; run += v[r]; }
}
__global__ void __launch_bounds__(256) k_cand_scatter(const uint2* __restrict__ cand, int64_t cand_cap,
                                                      const unsigned long long* __restrict__ counters, int shift,
                                                      uint32_t* __restrict__ offs, uint2* __restrict__ out) {
    const int64_t n_cand = min((int64_t)counters[1], cand_cap);
    for (int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_cand; x += (int64_t)gridDim.x * blockDim.x) {
        const uint2 c = __ldg(cand + x);
        out[atomicAdd(offs + (size_t)((c.x >> 1) >> shift) * CAND_PAD, 1u)] = c;
    }
}

// ---- phase 2: verify.  Thread per candidate (grid-stride).  A candidate survives when the two
// 16-mers are equal on the implied strand, the exact run through them is >= 28 long, and it is
// the leftmost position of that run indexed in BOTH samples.  Every window of 13 consecutive
// 16-mers lying inside a shared exact run has the same minimum in both samples, so a run that
// reaches 13 or more bases to the left of the anchor always holds an earlier joint position: such
// candidates are rejected after one 32-base compare, the others check the two index bitmaps over
// the (< 13) bases to the left.
// selected-in-both bits of the positions [p0, p0 + 32) of A (bit i <-> p0 + i) on diagonal d
__device__ __forceinline__ uint32_t joint_sel32(const PairView& P, int64_t ba, int64_t bb, int32_t lb, int32_t p0,
                                                int32_t d, int minus) {
    const uint32_t wa = fetchmask32(P.selA, ba + p0);
    uint32_t wb;
    if (!minus) wb = fetchmask32(P.selB, bb + p0 + d);
    else wb = __brev(fetchmask32(P.selB, bb + lb - d - RCQ_KMER - p0 - 31));
    return wa & wb;
}
// does [rs, re) hold RCQ_WORD consecutive positions with no seeding-mask bit in either sample?  (mB is the
// seeding mask of B on the strand of the hit, indexed like the bases: position p of A <-> p + d of B)
__device__ __forceinline__ bool clean_window(const uint32_t* __restrict__ mA, int64_t ba, const uint32_t* __restrict__ mB,
                                             int64_t bb, int32_t d, int32_t rs, int32_t re) {
    int32_t zr = 0;                       // clean positions ending at the current scan position
    for (int32_t p = rs; p < re; p += 32) {
        uint32_t w = fetchmask32(mA, ba + p) | fetchmask32(mB, bb + p + d);
        const int32_t nb = re - p;
        if (nb < 32) w |= ~0u << nb;
        if (w == 0u) { zr += 32; if (zr >= RCQ_WORD) return true; continue; }
        if (zr + (__ffs(w) - 1) >= RCQ_WORD) return true;
        uint32_t y = ~w;                  // a run of 28 clean positions inside this word
        y &= y >> 1; y &= y >> 2; y &= y >> 4; y &= y >> 8;
        if (y & (y >> 12)) return true;
        zr = __clz(w);
    }
    return false;
}

// A candidate becomes the seed of its exact run when it is the LEFTMOST position of the run indexed in both
// samples ("joint anchor") and the run qualifies: >= 28 bases with a 28-base window clean in both seeding
// masks.  Inside such a window all 13 16-mers are indexed candidates in both samples and share their minimum,
// so every qualifying run has a joint anchor.  Is there a joint anchor further left?  When the run reaches
// fewer than 13 bases to the left the two index bitmaps answer directly; when it reaches 13 or more and the
// 28 bases [ia - 13, ia + 15) are clean in both samples that window's minimum is an earlier joint anchor (the
// common case: one 32-base compare); only next to masked bases the bitmaps are scanned down the whole run.
__device__ __forceinline__ bool verify_anchor(const PairView& P, int32_t pA, int32_t pB, int minus, Seed* out) {
    const int32_t ta = __ldg(P.blkA + (pA >> 5)), tb = __ldg(P.blkB + (pB >> 5));
    const int2 a = __ldg(P.olA + ta), b = __ldg(P.olB + tb);
    const int32_t oa = a.x, la = a.y, ob = b.x, lb = b.y;
    const int32_t ia = pA - oa, jf = pB - ob;
    const int32_t j = minus ? lb - jf - RCQ_KMER : jf;
    const SeqArena& B = minus ? P.Br : P.Bf;
    const uint32_t* smB = minus ? P.smBr : P.smB;
    const int64_t ba = oa, bb = ob;
    if ((match32(P.Af, ba + ia, B, bb + j) & 0xFFFFu) != 0xFFFFu) return false;
    const int32_t d = j - ia;
    int32_t left = 0;
    const int32_t llim = min(ia, j);
    if (llim > 0) {
        uint32_t m = match32_back(P.Af, ba + ia - 1, B, bb + j - 1);
        left = (m == 0xFFFFFFFFu) ? 32 : (__ffs(~m) - 1);
        if (left > llim) left = llim;
        if (left >= RCQ_WIN) {
            const uint32_t W28 = (1u << RCQ_WORD) - 1u;
            if (!((fetchmask32(P.smA, ba + ia - RCQ_WIN) | fetchmask32(smB, bb + j - RCQ_WIN)) & W28)) return false;
            // masked bases nearby: walk the run to its start, looking for an earlier joint anchor
            int32_t pos = ia, rem = llim;
            left = 0;
            while (rem > 0) {
                m = match32_back(P.Af, ba + pos - 1, B, bb + pos + d - 1);
                int32_t z = (m == 0xFFFFFFFFu) ? 32 : (__ffs(~m) - 1);
                if (z > rem) z = rem;
                if (z > 0 && (joint_sel32(P, ba, bb, lb, pos - z, d, minus) & (z == 32 ? 0xFFFFFFFFu : ((1u << z) - 1u)))) return false;
                left += z; pos -= z; rem -= z;
                if (z < 32) break;
            }
        } else if (left > 0) {
            if (joint_sel32(P, ba, bb, lb, ia - left, d, minus) & ((1u << left) - 1u)) return false;
        }
    }
    const int32_t right = run_fwd(P.Af, ba + ia + RCQ_KMER, B, bb + j + RCQ_KMER,
                                  min(la - ia - RCQ_KMER, lb - j - RCQ_KMER));
    const int32_t rs = ia - left, re = ia + RCQ_KMER + right;
    if (re - rs < RCQ_WORD) return false;
    if (!clean_window(P.smA, ba, smB, bb, d, rs, re)) return false;
    out->ta = ta; out->tb = tb; out->minus = minus; out->rs = rs; out->re = re; out->diag = d;
    return true;
}

__global__ void __launch_bounds__(256) k_seed_verify(const uint2* __restrict__ cand, int64_t cand_cap, PairView P,
                                                     Seed* __restrict__ seeds, int64_t cap,
                                                     unsigned long long* __restrict__ counters /* [0]=seeds */) {
    const int lane = threadIdx.x & 31;
    // the candidate count is read from device memory: no host round trip between probe and verify
    const int64_t n_cand = min((int64_t)counters[1], cand_cap);
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < n_cand; base += (int64_t)gridDim.x * blockDim.x) {
        int64_t x = base + threadIdx.x;
        Seed s[2];
        int ns = 0;
        if (x < n_cand) {
            const uint2 c = __ldg(cand + x);
            const int32_t pA = (int32_t)(c.x >> 1), pB = (int32_t)(c.y >> 1);
            const int minus = (int)((c.x ^ c.y) & 1u);
            // palindromic 16-mers match on both strands
            uint32_t f = (uint32_t)fetch32(P.Af.bits, pA);
            uint32_t r = __brev(~f);
            r = ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
            const int reps = (r == f) ? 2 : 1;
            for (int rep = 0; rep < reps; rep++)
                if (verify_anchor(P, pA, pB, minus ^ rep, &s[ns])) ns++;
        }
        // warp-aggregated append
        unsigned m1 = __ballot_sync(0xffffffffu, ns >= 1), m2 = __ballot_sync(0xffffffffu, ns >= 2);
        int tot = __popc(m1) + __popc(m2);
        if (tot) {
            unsigned long long g = 0;
            if (lane == 0) g = atomicAdd(counters, (unsigned long long)tot);
            g = __shfl_sync(0xffffffffu, g, 0);
            unsigned below = (1u << lane) - 1u;
            if (ns >= 1) { int64_t slot = (int64_t)g + __popc(m1 & below); if (slot < cap) seeds[slot] = s[0]; }
            if (ns >= 2) { int64_t slot = (int64_t)g + __popc(m1) + __popc(m2 & below); if (slot < cap) seeds[slot] = s[1]; }
        }
    }
}

// X-drop ungapped extension of one seed + start point of the gapped extension.
// Word-parallel: 32 base pairs per match mask; a run of matches is consumed in one step.
__device__ __forceinline__ Ungapped ungapped_one(const Seed& s, const PairView& P, unsigned long long& work) {
    Ungapped u;
    {

        const SeqArena& B = s.minus ? P.Br : P.Bf;
        int64_t ba = __ldg(P.offA + s.ta), bb = __ldg(P.offB + s.tb);
        int32_t la = __ldg(P.lenA + s.ta), lb = __ldg(P.lenB + s.tb), d = s.diag;
        // ---- right
        int32_t sum = 0, best = 0, ext = 0, k = 0, lim = min(la - s.re, lb - s.re - d);
        bool stop = false;
        while (k < lim && !stop) {
            uint32_t m = match32(P.Af, ba + s.re + k, B, bb + s.re + d + k);
            int32_t v = min(32, lim - k), i = 0;
            while (i < v) {
                uint32_t rest = ~(m >> i);
                int32_t z = rest ? (__ffs(rest) - 1) : 32;
                if (z > v - i) z = v - i;
                if (z) { sum += z; i += z; if (sum > best) { best = sum; ext = k + i; } }
                if (i < v) {  // mismatch at bit i
                    sum -= 2; i++;
                    if (best - sum > RCQ_XDROP_UNGAPPED) { stop = true; break; }
                }
            }
            k += i;
        }
        work += k;
        int32_t a1 = s.re + ext, score = (s.re - s.rs) + best;
        // ---- left
        sum = 0; best = 0; ext = 0; k = 0; lim = min(s.rs, s.rs + d); stop = false;
        while (k < lim && !stop) {
            uint32_t m = match32_back(P.Af, ba + s.rs - 1 - k, B, bb + s.rs - 1 + d - k);
            int32_t v = min(32, lim - k), i = 0;
            while (i < v) {
                uint32_t rest = ~(m >> i);
                int32_t z = rest ? (__ffs(rest) - 1) : 32;
                if (z > v - i) z = v - i;
                if (z) { sum += z; i += z; if (sum > best) { best = sum; ext = k + i; } }
                if (i < v) {
                    sum -= 2; i++;
                    if (best - sum > RCQ_XDROP_UNGAPPED) { stop = true; break; }
                }
            }
            k += i;
        }
        work += k;
        int32_t a0 = s.rs - ext;
        score += best;
        // ---- longest exact run inside [a0, a1) (earliest on ties) -> gapped start point
        int32_t run = 0, best_len = 0, best_s = a0;
        for (int32_t p = a0; p < a1; p += 32) {
            uint32_t m = match32(P.Af, ba + p, B, bb + p + d);
            int32_t v = min(32, a1 - p), i = 0;
            while (i < v) {
                uint32_t rest = ~(m >> i);
                int32_t z = rest ? (__ffs(rest) - 1) : 32;
                if (z > v - i) z = v - i;
                run += z; i += z;
                if (i < v) {
                    if (run > best_len) { best_len = run; best_s = p + i - run; }
                    run = 0; i++;
                }
            }
        }
        if (run > best_len) { best_len = run; best_s = a1 - run; }
        work += a1 - a0;
        u.ta = s.ta; u.tb = s.tb; u.minus = s.minus; u.diag = d; u.a0 = a0; u.a1 = a1;
        u.score = score; u.i0 = best_s + best_len / 2;
    }
    return u;
}

// sort key of a seed: (ta, tb, strand, diagonal); seeds of one diagonal end up adjacent, and so do
// the diagonals of one (ta, tb, strand) group
// The key is packed to the bit widths this pair needs (KeyFmt), so the radix sort runs over
// ~46 bits instead of 64: [ta | tb (tb_bits) | strand (1) | diag + bias (d_bits)].
struct KeyFmt { int32_t d_bits, tb_bits, bias, total_bits; };
__global__ void k_seed_keys(const Seed* __restrict__ seeds, int64_t n, KeyFmt F, uint64_t* __restrict__ keys,
                            uint32_t* __restrict__ vals) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const Seed s = seeds[e];
    keys[e] = ((((((uint64_t)(uint32_t)s.ta << F.tb_bits) | (uint64_t)(uint32_t)s.tb) << 1) | (uint64_t)s.minus) << F.d_bits) |
              (uint64_t)(uint32_t)(s.diag + F.bias);
    vals[e] = (uint32_t)e;
}
// heads of diagonals and of (ta, tb, strand) groups in the sorted order (list order is irrelevant)
__global__ void k_seed_heads(const uint64_t* __restrict__ keys, int64_t n, KeyFmt F, uint32_t* __restrict__ dheads,
                             uint32_t* __restrict__ gheads, unsigned long long* __restrict__ counters /* [5]=groups [6]=diagonals */) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const uint64_t k = keys[e], kp = e ? keys[e - 1] : ~k;
    if (k != kp) dheads[atomicAdd(counters + 6, 1ull)] = (uint32_t)e;
    if ((k >> F.d_bits) != (kp >> F.d_bits)) gheads[atomicAdd(counters + 5, 1ull)] = (uint32_t)e;
}
// BLAST's diagonal rule, thread per diagonal: seeds in ascending position; a seed that starts inside
// the ungapped HSP extended from an earlier seed of the diagonal is skipped.  U is indexed like the
// sorted seed list; slots of skipped seeds keep score = -1.
__global__ void __launch_bounds__(128) k_ungapped_diag(const Seed* __restrict__ seeds, const uint32_t* __restrict__ order,
                                                       const uint64_t* __restrict__ keys, int64_t n,
                                                       const uint32_t* __restrict__ dheads, int64_t n_diag, PairView P,
                                                       Ungapped* __restrict__ U,
                                                       unsigned long long* __restrict__ counters /* [4]=bases [7]=HSPs */) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long work = 0, made = 0;
    if (g < n_diag) {
        const int64_t h = dheads[g];
        const uint64_t key = keys[h];
        int64_t e = h + 1;
        while (e < n && keys[e] == key) e++;
        int32_t last_end = -1;
        while (true) {
            int64_t pick = -1;
            int32_t pick_rs = 0x7fffffff;
            for (int64_t x = h; x < e; x++) {
                int32_t rs = seeds[order[x]].rs;
                if (rs >= last_end && rs < pick_rs) { pick_rs = rs; pick = x; }
            }
            if (pick < 0) break;
            const Seed s = seeds[order[pick]];
            Ungapped u = ungapped_one(s, P, work);
            U[pick] = u;
            made++;
            last_end = max(u.a1, s.rs + 1);
        }
    }
    for (int o = 16; o; o >>= 1) {
        work += __shfl_down_sync(0xffffffffu, work, o);
        made += __shfl_down_sync(0xffffffffu, made, o);
    }
    if ((threadIdx.x & 31) == 0 && work) { atomicAdd(counters + 4, work); atomicAdd(counters + 7, made); }
}


// ---- diagonal table: the first ungapped HSP of every diagonal without sorting the seeds ----------
// Between orthologs ~95 % of the seeds lie inside the ungapped HSP grown from the leftmost seed of
// their diagonal and are skipped by the diagonal rule.  So: (1) every seed enters an open-addressing
// table keyed by its packed (ta, tb, strand, diagonal), the slot keeps the leftmost seed
// (atomicMin on rs<<32 | seed index); (2) one thread per occupied slot extends that seed; (3) every
// seed compares its start with the end of its diagonal's first HSP — the few that start beyond it
// ("survivors") go through the sort + k_ungapped_diag path, which continues the diagonal rule exactly
// where step (2) left it (all of them start at or after that HSP's end).
#define DIAG_EMPTY 0xFFFFFFFFFFFFFFFFull
struct DiagTable { unsigned long long* keys; unsigned long long* vals; uint32_t mask; };

__device__ __forceinline__ uint64_t seed_key(const Seed& s, const KeyFmt& F) {
    return ((((((uint64_t)(uint32_t)s.ta << F.tb_bits) | (uint64_t)(uint32_t)s.tb) << 1) | (uint64_t)s.minus) << F.d_bits) |
           (uint64_t)(uint32_t)(s.diag + F.bias);
}
__device__ __forceinline__ uint32_t diag_hash(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 29;
    return (uint32_t)k;
}

__global__ void k_diag_insert(const Seed* __restrict__ seeds, int64_t n, KeyFmt F, DiagTable T,
                              uint32_t* __restrict__ slots, unsigned long long* __restrict__ counters /* [6]=diagonals */) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const Seed s = seeds[e];
    const unsigned long long key = seed_key(s, F);
    uint32_t h = diag_hash(key) & T.mask;
    while (true) {
        const unsigned long long old = atomicCAS(T.keys + h, DIAG_EMPTY, key);
        if (old == DIAG_EMPTY) { slots[atomicAdd(counters + 6, 1ull)] = h; break; }
        if (old == key) break;
        h = (h + 1) & T.mask;
    }
    atomicMin(T.vals + h, ((unsigned long long)(uint32_t)s.rs << 32) | (unsigned long long)(uint32_t)e);
}

// X-drop extension going forward from (pa, pb): best gain and the bases it spans (the right half of
// ungapped_one; the left half is the same walk on the reverse-complement arenas)
__device__ __forceinline__ void xdrop_fwd(const SeqArena& A, int64_t pa, const SeqArena& B, int64_t pb, int32_t lim,
                                          int32_t& best, int32_t& ext, unsigned long long& work) {
    int32_t sum = 0, k = 0;
    best = 0; ext = 0;
    bool stop = false;
    while (k < lim && !stop) {
        const uint32_t m = match32(A, pa + k, B, pb + k);
        const int32_t v = min(32, lim - k);
        int32_t i = 0;
        while (i < v) {
            const uint32_t rest = ~(m >> i);
            int32_t z = rest ? (__ffs(rest) - 1) : 32;
            if (z > v - i) z = v - i;
            if (z) { sum += z; i += z; if (sum > best) { best = sum; ext = k + i; } }
            if (i < v) {  // mismatch at bit i
                sum -= 2; i++;
                if (best - sum > RCQ_XDROP_UNGAPPED) { stop = true; break; }
            }
        }
        k += i;
    }
    work += k;
}
// longest exact run inside [p0, p1) of diagonal d (earliest on ties), plus the runs touching both ends
struct RunScan { int32_t best_len, best_s, lead, tail; };
__device__ __forceinline__ RunScan run_scan(const SeqArena& A, int64_t ba, const SeqArena& B, int64_t bbd,
                                            int32_t p0, int32_t p1) {
    RunScan r;
    int32_t run = 0;
    r.best_len = 0; r.best_s = p0; r.lead = -1;
    for (int32_t p = p0; p < p1; p += 32) {
        const uint32_t m = match32(A, ba + p, B, bbd + p);
        const int32_t v = min(32, p1 - p);
        int32_t i = 0;
        while (i < v) {
            const uint32_t rest = ~(m >> i);
            int32_t z = rest ? (__ffs(rest) - 1) : 32;
            if (z > v - i) z = v - i;
            run += z; i += z;
            if (i < v) {
                if (r.lead < 0) r.lead = run;
                if (run > r.best_len) { r.best_len = run; r.best_s = p + i - run; }
                run = 0; i++;
            }
        }
    }
    if (run > r.best_len) { r.best_len = run; r.best_s = p1 - run; }
    if (r.lead < 0) r.lead = run;
    r.tail = run;
    return r;
}

// Two lanes per occupied slot (count read on the device): the even lane extends the leftmost seed of
// the diagonal to the right, the odd lane to the left — as a forward walk on the reverse-complement
// arenas, so both lanes run the same code — then each scans half of the HSP for its longest exact run.
__global__ void __launch_bounds__(128) k_diag_first(const Seed* __restrict__ seeds, const uint32_t* __restrict__ slots,
                                                    DiagTable T, PairView P, Ungapped* __restrict__ U,
                                                    unsigned long long* __restrict__ counters /* [4]=bases [6]=diagonals */) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 1;
    const int role = threadIdx.x & 1;
    unsigned long long work = 0;
    const bool on = i < (int64_t)counters[6];
    uint32_t h = 0;
    Seed s;
    s.ta = s.tb = s.minus = s.rs = s.re = s.diag = 0;
    int32_t best = 0, ext = 0, la = 0, lb = 0;
    int64_t ba = 0, bb = 0;
    if (on) {
        h = slots[i];
        s = seeds[(uint32_t)T.vals[h]];
        ba = __ldg(P.offA + s.ta); bb = __ldg(P.offB + s.tb);
        la = __ldg(P.lenA + s.ta); lb = __ldg(P.lenB + s.tb);
        const int32_t d = s.diag;
        if (role == 0) {
            const SeqArena& B = s.minus ? P.Br : P.Bf;
            xdrop_fwd(P.Af, ba + s.re, B, bb + s.re + d, min(la - s.re, lb - s.re - d), best, ext, work);
        } else {
            // bases rs-1, rs-2, ... of A and rs-1+d, ... of B, read forward on the opposite strands
            const SeqArena& Bo = s.minus ? P.Bf : P.Br;
            xdrop_fwd(P.Ar, ba + (la - s.rs), Bo, bb + (lb - s.rs - d), min(s.rs, s.rs + d), best, ext, work);
        }
    }
    const int32_t obest = __shfl_xor_sync(0xffffffffu, best, 1), oext = __shfl_xor_sync(0xffffffffu, ext, 1);
    const int32_t a1 = s.re + (role == 0 ? ext : oext), a0 = s.rs - (role == 0 ? oext : ext);
    const int32_t score = (s.re - s.rs) + best + obest;
    const int32_t mid = a0 + ((a1 - a0) >> 1);
    RunScan r;
    r.best_len = 0; r.best_s = 0; r.lead = 0; r.tail = 0;
    if (on) {
        const SeqArena& B = s.minus ? P.Br : P.Bf;
        r = run_scan(P.Af, ba, B, bb + s.diag, role == 0 ? a0 : mid, role == 0 ? mid : a1);
        work += (role == 0 ? mid - a0 : a1 - mid);
    }
    RunScan y;
    y.best_len = __shfl_xor_sync(0xffffffffu, r.best_len, 1); y.best_s = __shfl_xor_sync(0xffffffffu, r.best_s, 1);
    y.lead = __shfl_xor_sync(0xffffffffu, r.lead, 1); y.tail = __shfl_xor_sync(0xffffffffu, r.tail, 1);
    if (on && role == 0) {
        int32_t bl = r.best_len, bs = r.best_s;                  // left half, then the run across the middle, then the right half
        if (r.tail + y.lead > bl) { bl = r.tail + y.lead; bs = mid - r.tail; }
        if (y.best_len > bl) { bl = y.best_len; bs = y.best_s; }
        Ungapped u;
        u.ta = s.ta; u.tb = s.tb; u.minus = s.minus; u.diag = s.diag; u.a0 = a0; u.a1 = a1;
        u.score = score; u.i0 = bs + bl / 2;
        U[i] = u;
        T.vals[h] = (unsigned long long)(uint32_t)max(a1, s.rs + 1);   // end of the diagonal's first HSP
    }
    for (int o = 16; o; o >>= 1) work += __shfl_down_sync(0xffffffffu, work, o);
    if ((threadIdx.x & 31) == 0 && work) atomicAdd(counters + 4, work);
}

__global__ void k_diag_survivors(const Seed* __restrict__ seeds, int64_t n, KeyFmt F, DiagTable T,
                                 Seed* __restrict__ surv, unsigned long long* __restrict__ counters /* [9]=survivors */) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const Seed s = seeds[e];
    const unsigned long long key = seed_key(s, F);
    uint32_t h = diag_hash(key) & T.mask;
    while (T.keys[h] != key) h = (h + 1) & T.mask;
    if (s.rs >= (int32_t)(uint32_t)T.vals[h]) surv[atomicAdd(counters + 9, 1ull)] = s;
}

// ---- ungapped HSPs grouped by (ta, tb, strand) for the gapped kernel -----------------------------
// group key = seed key without the diagonal; an empty slot (score < 0) sorts behind every group
__global__ void k_hsp_keys(const Ungapped* __restrict__ U, int64_t n, KeyFmt F, uint64_t* __restrict__ keys,
                           uint32_t* __restrict__ vals, unsigned long long* __restrict__ counters /* [7]=valid HSPs */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = false;
    if (i < n) {
        const Ungapped u = U[i];
        valid = u.score >= 0;
        const int gbits = F.total_bits - F.d_bits;
        keys[i] = valid ? (((((uint64_t)(uint32_t)u.ta << F.tb_bits) | (uint64_t)(uint32_t)u.tb) << 1) | (uint64_t)u.minus)
                        : (1ull << gbits);
        vals[i] = (uint32_t)i;
    }
    const unsigned m = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(counters + 7, (unsigned long long)__popc(m));
}
__global__ void k_hsp_gather(const Ungapped* __restrict__ U, const uint32_t* __restrict__ order,
                             const uint64_t* __restrict__ keys, int64_t n, KeyFmt F, Ungapped* __restrict__ Us,
                             uint32_t* __restrict__ gheads, unsigned long long* __restrict__ counters /* [5]=groups */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Us[i] = U[order[i]];
    const uint64_t k = keys[i];
    if ((k >> (F.total_bits - F.d_bits)) == 0 && (i == 0 || keys[i - 1] != k)) gheads[atomicAdd(counters + 5, 1ull)] = (uint32_t)i;
}
