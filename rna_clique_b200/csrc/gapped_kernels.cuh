// gapped_kernels.cuh — kernel (3b): greedy X-drop gapped extension (megablast's own algorithm for
// linear gap costs: Zhang, Schwartz, Wagner, Miller 2000; NCBI BLAST_GreedyAlign), scoring +1/-2,
// gap 2.5 per position.  In doubled units a path that ends on anti-diagonal t = i + j after d
// differences (mismatches + gap characters) scores exactly t - 6d, so per difference count d and
// diagonal k = j - i only the furthest row R(d, k) is kept ("wave" d).  X-drop 100 bits = 54 raw
// = 108 doubled, applied BLAST's way: an entry of wave d is dropped when its score is more than
// 108 below the best score of waves <= d - 19 (19 = floor((108 + 1) / 6) + 1).
//
// One warp per (transcript of A, transcript of B, strand) group.  A wave is processed 32 diagonals
// at a time, one per lane: the three candidate entries come from the previous wave's row (a
// shared-memory ring of 128 diagonals per warp; global scratch rows when a wave is wider), the
// slide over matching bases is word-parallel (32 bases per XOR/compare of 2-bit words), and a
// slide that survives its first 32 bases is finished by the whole warp together (1024 bases per
// round).  Mismatch / gap / gap-opening counts travel with every (d, k) point, so there is no
// traceback pass.  The wave loop stops early once no live diagonal can still beat the best score
// (upper bound 2*min(qn, sn-k) + k - 6d), which ends a full-length extension the moment it
// reaches the transcript ends.
//
// Roofline: instruction-issue bound (L2-resident 2-bit words; no HBM traffic to speak of).
// Work unit: base comparisons ("extension steps"), reported per search.
#pragma once
#include "search.cuh"

#define NEG (-(1 << 29))
#define FULL 0xffffffffu
#define RCQ_RING 128            // diagonals per shared-memory row
#define RCQ_XDROP_OFFSET 19     // floor((108 + 1) / 6) + 1
#define RCQ_STAGE_MAX 4096      // longest transcript staged in shared memory by the fast path
#define RCQ_STAGE_W (RCQ_STAGE_MAX / 16 + 4)

struct ExtRes { int32_t score2, qi, sj, mism, gw; };  // gw = gaps<<17 | gapopen<<3 | slid<<2 | entry op

// number of matching base pairs from (qp, sp), at most lim (lim >= 0), computed by the whole warp
__device__ __forceinline__ int32_t coop_slide(const SeqArena& Q, int64_t qp, const SeqArena& S, int64_t sp,
                                              int32_t lim) {
    const int lane = threadIdx.x & 31;
    int32_t base = 0;
    while (base < lim) {
        int32_t o = base + 32 * lane, z = 32;
        if (o < lim) {
            uint32_t m = match32(Q, qp + o, S, sp + o);
            if (m != 0xFFFFFFFFu) z = __ffs(~m) - 1;
            if (z > lim - o) z = lim - o;
        } else z = 0;
        unsigned stop = __ballot_sync(FULL, z < 32);
        if (stop) {
            int src = __ffs(stop) - 1;
            return base + 32 * src + __shfl_sync(FULL, z, src);
        }
        base += 1024;
    }
    return lim;
}

// One extension by one warp.  rows: 6 arrays of rowlen int32 (R[2], M[2], G[2]); RING: index
// k & (RCQ_RING-1) (returns false when a wave is wider than the ring), else index k + qn + 1.
template <bool RING>
__device__ bool greedy_extend(const SeqArena Q, int64_t qpos, int32_t qn, const SeqArena S, int64_t spos,
                              int32_t sn, int32_t* rows, int32_t rowlen, ExtRes* res,
                              unsigned long long* steps) {
    const int lane = threadIdx.x & 31;
    int32_t* Rr = rows; int32_t* Mr = rows + 2 * rowlen; int32_t* Gr = rows + 4 * rowlen;
#define IX(k) (RING ? ((k) & (RCQ_RING - 1)) : ((k) + qn + 1))
    // wave 0
    int32_t i0 = coop_slide(Q, qpos, S, spos, min(qn, sn));
    unsigned long long nstep = (lane == 0) ? (unsigned long long)(i0 + 1) : 0ull;
    if (lane == 0) { Rr[IX(0)] = i0; Mr[IX(0)] = 0; Gr[IX(0)] = 0; }
    int32_t bS = (lane == 0) ? 2 * i0 : NEG, bD = 0, bK = 0, bI = i0, bM = 0, bG = 0;
    int32_t gb = 2 * i0;                 // best score so far (uniform)
    int32_t Tl = (lane == 0) ? gb : 0;   // lane (d & 31) holds T[d]
    int32_t Tprev = gb, L = 0, U = 0;
    bool ok = true;
    __syncwarp();
    for (int32_t d = 1; d <= qn + sn; d++) {
        const int32_t p = ((d - 1) & 1) * rowlen, c = (d & 1) * rowlen;
        int32_t tref = __shfl_sync(FULL, Tl, (d - RCQ_XDROP_OFFSET) & 31);
        if (d < RCQ_XDROP_OFFSET) tref = 0;
        const int32_t klo = max(L - 1, -qn), khi = min(U + 1, sn);
        if (RING && khi - klo + 1 > RCQ_RING) { ok = false; break; }
        int32_t nL = 1 << 28, nU = -(1 << 28), wbest = NEG;
        for (int32_t kb = klo; kb <= khi; kb += 32) {
            const int32_t k = kb + lane;
            int32_t bi = -1, src = 0, op = 0;
            if (k <= khi) {
                if (k >= L && k <= U) {
                    int32_t r = Rr[p + IX(k)];
                    if (r >= 0 && r + 1 <= qn && r + 1 + k <= sn) { bi = r + 1; src = k; op = 0; }
                }
                if (k + 1 >= L && k + 1 <= U) {
                    int32_t r = Rr[p + IX(k + 1)];
                    if (r >= 0 && r + 1 <= qn && r + 1 > bi) { bi = r + 1; src = k + 1; op = 1; }
                }
                if (k - 1 >= L && k - 1 <= U) {
                    int32_t r = Rr[p + IX(k - 1)];
                    if (r >= 0 && r + k <= sn && r > bi) { bi = r; src = k - 1; op = 2; }
                }
                if (bi >= 0 && 2 * bi + k - 6 * d < tref - RCQ_XDROP_GAPPED2) bi = -1;
            }
            int32_t e = bi, mm = 0, gw = 0, lim = 0;
            bool cont = false;
            if (bi >= 0) {
                mm = Mr[p + IX(src)];
                int32_t g = Gr[p + IX(src)];
                if (op == 0) { mm++; gw = g & ~7; }
                else gw = ((g & ~7) | op) + (1 << 17) + (((g & 7) != op) ? 8 : 0);
                lim = min(qn - bi, sn - bi - k);
                int32_t z = 0;
                if (lim > 0) {
                    uint32_t m = match32(Q, qpos + bi, S, spos + bi + k);
                    z = (m == 0xFFFFFFFFu) ? 32 : (__ffs(~m) - 1);
                    if (z > lim) z = lim;
                }
                nstep += z + 1;
                e = bi + z;
                cont = (z == 32 && lim > 32);
            }
            unsigned more = __ballot_sync(FULL, cont);
            while (more) {
                int sl = __ffs(more) - 1;
                more &= more - 1;
                int32_t e2 = __shfl_sync(FULL, e, sl), k2 = __shfl_sync(FULL, k, sl), l2 = __shfl_sync(FULL, lim, sl);
                int32_t extra = coop_slide(Q, qpos + e2, S, spos + e2 + k2, l2 - 32);
                if (lane == sl) { e += extra; nstep += extra; }
            }
            if (k <= khi) {
                if (bi >= 0) {
                    if (e > bi) gw |= 4;
                    Rr[c + IX(k)] = e; Mr[c + IX(k)] = mm; Gr[c + IX(k)] = gw;
                    int32_t sc = 2 * e + k - 6 * d;
                    wbest = max(wbest, sc);
                    nL = min(nL, k); nU = max(nU, k);
                    if (sc > bS) { bS = sc; bD = d; bK = k; bI = e; bM = mm; bG = gw; }
                } else {
                    Rr[c + IX(k)] = -1;
                }
            }
        }
        wbest = __reduce_max_sync(FULL, wbest);
        nL = __reduce_min_sync(FULL, nL);
        nU = __reduce_max_sync(FULL, nU);
        Tprev = max(Tprev, wbest);
        if (lane == (d & 31)) Tl = Tprev;
        gb = max(gb, wbest);
        L = nL; U = nU;
        __syncwarp();
        if (L > U) break;
        {   // no diagonal of the band can still beat the best end point, even by matching to the end of q or s
            const int32_t kp = min(max(sn - qn, L), U);
            if (2 * min(qn, sn - kp) + kp - 6 * d <= gb) break;
        }
    }
#undef IX
    for (int o = 16; o; o >>= 1) nstep += __shfl_down_sync(FULL, nstep, o);
    if (lane == 0) *steps += nstep;
    if (!ok) return false;
    // best end point: score desc, then d asc, then k asc
    uint64_t key = 0;
    if (bS > NEG)
        key = ((uint64_t)(uint32_t)(bS + (1 << 20)) << 42) | ((uint64_t)((1 << 21) - 1 - bD) << 21) |
              (uint64_t)((1 << 21) - 1 - (bK + (1 << 20)));
    uint64_t best = key;
    for (int o = 16; o; o >>= 1) { uint64_t x = __shfl_xor_sync(FULL, best, o); if (x > best) best = x; }
    int src = __ffs(__ballot_sync(FULL, key == best)) - 1;
    res->score2 = __shfl_sync(FULL, bS, src);
    res->qi = __shfl_sync(FULL, bI, src);
    res->sj = res->qi + __shfl_sync(FULL, bK, src);
    res->mism = __shfl_sync(FULL, bM, src);
    res->gw = __shfl_sync(FULL, bG, src);
    return true;
}

// ---- fast path for samples without masked bases (no N): same recurrence, fewer instructions ----
// 16 bases per first slide step from two 32-bit words (one funnel shift, no 64-bit compress), 32-bit
// offsets from per-extension base pointers, dead cells stored as NEG with two sentinels either side
// of every wave so the three-way max needs no range checks.
// Sequence accessors of the fast path: words of one strand of one transcript, either straight from
// the arena (read-only cache) or from the warp's staged copy in shared memory (LDS, 32-bit addresses).
struct SeqGlobal {
    const uint32_t* b;
    __device__ __forceinline__ uint32_t w(int32_t i) const { return __ldg(b + i); }
};
struct SeqShared {
    const uint32_t* b;
    __device__ __forceinline__ uint32_t w(int32_t i) const { return b[i]; }
};
template <typename SEQ>
__device__ __forceinline__ uint32_t fetch16(const SEQ& q, int32_t x) {
    const int32_t w = x >> 4;
    return __funnelshift_r(q.w(w), q.w(w + 1), (x & 15) * 2);
}
template <typename SEQ>
__device__ __forceinline__ int32_t run16(const SEQ& Q, int32_t xq, const SEQ& S, int32_t xs) {
    uint32_t d = fetch16(Q, xq) ^ fetch16(S, xs);
    d = (d | (d >> 1)) & 0x55555555u;
    return d ? ((__ffs(d) - 1) >> 1) : 16;
}
template <typename SEQ>
__device__ __forceinline__ int32_t coop_slide16(const SEQ& Q, int32_t xq, const SEQ& S, int32_t xs, int32_t lim) {
    const int lane = threadIdx.x & 31;
    int32_t base = 0;
    while (base < lim) {
        const int32_t o = base + 16 * lane;
        int32_t z = 0;
        if (o < lim) z = min(run16(Q, xq + o, S, xs + o), lim - o);
        const unsigned stop = __ballot_sync(FULL, z < 16);
        if (stop) {
            const int src = __ffs(stop) - 1;
            return base + 16 * src + __shfl_sync(FULL, z, src);
        }
        base += 512;
    }
    return lim;
}

// rows: R (2 x rowlen int32) followed by MG (2 x rowlen uint64: mismatches | gap word << 32).
// RING: index k & (RCQ_RING-1); else index k + qn + 2 (rowlen >= qn+sn+5).
template <bool RING, typename SEQ>
__device__ bool greedy_fast(const SEQ Q, int32_t qo, int32_t qn, const SEQ S, int32_t so, int32_t sn, int32_t* rows,
                            int32_t rowlen, ExtRes* res, unsigned long long* steps) {
    const int lane = threadIdx.x & 31;
    int32_t* Rr = rows;
    uint2* MGr = (uint2*)(rows + 2 * rowlen);
#define IX(k) (RING ? ((k) + RCQ_RING / 2) : ((k) + qn + 2))   // RING: diagonals -62 .. 61 in place, no wrap
    const int32_t i0 = coop_slide16(Q, qo, S, so, min(qn, sn));
    uint32_t nstep = (lane == 0) ? (uint32_t)(i0 + 1) : 0u;
    if (lane == 0) { Rr[IX(0)] = i0; MGr[IX(0)] = make_uint2(0u, 0u); }
    if (lane >= 1 && lane <= 4) Rr[IX(lane <= 2 ? -lane : lane - 2)] = NEG;
    int32_t bS = (lane == 0) ? 2 * i0 : NEG, bD = 0, bK = 0, bI = i0, bM = 0, bG = 0;
    int32_t gb = 2 * i0;
    int32_t Tl = (lane == 0) ? gb : 0;   // lane (d & 31) holds T[d]
    int32_t Tprev = gb, L = 0, U = 0;
    bool ok = true;
    __syncwarp();
    for (int32_t d = 1; d <= qn + sn; d++) {
        const int32_t p = ((d - 1) & 1) * rowlen, c = (d & 1) * rowlen;
        int32_t tref = __shfl_sync(FULL, Tl, (d - RCQ_XDROP_OFFSET) & 31);
        if (d < RCQ_XDROP_OFFSET) tref = 0;
        const int32_t klo = max(L - 1, -qn), khi = min(U + 1, sn);
        if (RING && (klo - 2 < -RCQ_RING / 2 || khi + 2 >= RCQ_RING / 2)) { ok = false; break; }
        const int32_t xcut = tref - RCQ_XDROP_GAPPED2 + 6 * d;   // alive iff 2*row + k >= xcut
        const int32_t d6 = 6 * d;
        int32_t nL = 1 << 28, nU = -(1 << 28), wbest = NEG;
        for (int32_t kb = klo; kb <= khi; kb += 32) {
            const int32_t k = kb + lane;
            const bool active = k <= khi;
            const int32_t ix = IX(k);
            int32_t bi = NEG, op = 0, rowmax = 0, e = NEG, lim = 0;
            uint32_t mm = 0, gw = 0;
            bool cont = false;
            if (active) {
                rowmax = min(qn, sn - k);
                int32_t c0 = Rr[p + ix] + 1, c1 = Rr[p + IX(k + 1)] + 1, c2 = Rr[p + IX(k - 1)];
                if (c0 > rowmax) c0 = NEG;
                if (c1 > rowmax) c1 = NEG;
                if (c2 > rowmax) c2 = NEG;
                bi = c0;
                if (c1 > bi) { bi = c1; op = 1; }
                if (c2 > bi) { bi = c2; op = 2; }
                if (2 * bi + k < xcut) bi = NEG;
            }
            if (bi >= 0) {
                const int32_t src = k + (op == 1) - (op == 2);
                const uint2 mg = MGr[p + IX(src)];
                mm = mg.x;
                const uint32_t g = mg.y;
                if (op == 0) { mm++; gw = g & ~7u; }
                else gw = ((g & ~7u) | (uint32_t)op) + (1u << 17) + (((g & 7u) != (uint32_t)op) ? 8u : 0u);
                lim = rowmax - bi;
                int32_t z = 0;
                if (lim > 0) z = min(run16(Q, qo + bi, S, so + bi + k), lim);
                nstep += z + 1;
                e = bi + z;
                cont = (z == 16 && lim > 16);
            }
            unsigned more = __ballot_sync(FULL, cont);
            while (more) {
                const int sl = __ffs(more) - 1;
                more &= more - 1;
                const int32_t e2 = __shfl_sync(FULL, e, sl), k2 = __shfl_sync(FULL, k, sl), l2 = __shfl_sync(FULL, lim, sl);
                const int32_t extra = coop_slide16(Q, qo + e2, S, so + e2 + k2, l2 - 16);
                if (lane == sl) { e += extra; nstep += extra; }
            }
            if (active) {
                if (bi >= 0) {
                    if (e > bi) gw |= 4u;
                    Rr[c + ix] = e;
                    MGr[c + ix] = make_uint2(mm, gw);
                    const int32_t sc = 2 * e + k - d6;
                    wbest = max(wbest, sc);
                    nL = min(nL, k); nU = max(nU, k);
                    if (sc > bS) { bS = sc; bD = d; bK = k; bI = e; bM = (int32_t)mm; bG = (int32_t)gw; }
                } else {
                    Rr[c + ix] = NEG;
                }
            }
        }
        if (lane < 4) Rr[c + IX(lane < 2 ? klo - 1 - lane : khi + lane - 1)] = NEG;   // sentinels
        wbest = __reduce_max_sync(FULL, wbest);
        nL = __reduce_min_sync(FULL, nL);
        nU = __reduce_max_sync(FULL, nU);
        Tprev = max(Tprev, wbest);
        if (lane == (d & 31)) Tl = Tprev;
        gb = max(gb, wbest);
        L = nL; U = nU;
        __syncwarp();
        if (L > U) break;
        {   // no diagonal of the band can still beat the best end point, even by matching to the end of q or s
            const int32_t kp = min(max(sn - qn, L), U);
            if (2 * min(qn, sn - kp) + kp - d6 <= gb) break;
        }
    }
#undef IX
    nstep = __reduce_add_sync(FULL, nstep);
    if (lane == 0) *steps += nstep;
    if (!ok) return false;
    uint64_t key = 0;
    if (bS > NEG)
        key = ((uint64_t)(uint32_t)(bS + (1 << 20)) << 42) | ((uint64_t)((1 << 21) - 1 - bD) << 21) |
              (uint64_t)((1 << 21) - 1 - (bK + (1 << 20)));
    uint64_t best = key;
    for (int o = 16; o; o >>= 1) { uint64_t x = __shfl_xor_sync(FULL, best, o); if (x > best) best = x; }
    const int src = __ffs(__ballot_sync(FULL, key == best)) - 1;
    res->score2 = __shfl_sync(FULL, bS, src);
    res->qi = __shfl_sync(FULL, bI, src);
    res->sj = res->qi + __shfl_sync(FULL, bK, src);
    res->mism = __shfl_sync(FULL, bM, src);
    res->gw = __shfl_sync(FULL, bG, src);
    return true;
}

// ---- one extension: shared-memory ring first; a wave that outgrows the ring re-runs the extension
// on global rows.  Those rows come from a small pool of slots shared by all warps (try-lock per
// slot; a holder never waits on anything, so the spin cannot deadlock) instead of one private
// region per resident warp, which cost 0.5 GB per search for a path almost never taken.
struct ScratchPool { int32_t* rows; int64_t per_slot; unsigned int* locks; int32_t n_slots; };

__device__ __forceinline__ int32_t scratch_acquire(const ScratchPool& sp) {
    int32_t s = 0;
    if ((threadIdx.x & 31) == 0) {
        s = (int32_t)(((blockIdx.x * blockDim.x + threadIdx.x) >> 5) % sp.n_slots);
        while (atomicCAS(sp.locks + s, 0u, 1u) != 0u) s = (s + 1 == sp.n_slots) ? 0 : s + 1;
    }
    return __shfl_sync(FULL, s, 0);
}
__device__ __forceinline__ void scratch_release(const ScratchPool& sp, int32_t s) {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) { __threadfence(); atomicExch(sp.locks + s, 0u); }
}

// Fallback of the fast path on global rows (rare: a wave wider than the ring); kept out of line so the
// hot loop stays small in the instruction cache.
template <typename SEQ>
__device__ __noinline__ void greedy_fast_wide(const SEQ Q, int32_t qo, int32_t qn, const SEQ S, int32_t so, int32_t sn,
                                              const ScratchPool sp, ExtRes* r, unsigned long long* steps) {
    const int32_t slot = scratch_acquire(sp);
    greedy_fast<false, SEQ>(Q, qo, qn, S, so, sn, sp.rows + slot * sp.per_slot, qn + sn + 5, r, steps);
    scratch_release(sp, slot);
}

// One extension.  MASKED: general path on the arenas (mask words consulted).  Otherwise the fast path,
// on the warp's staged copies (stQ/stS != nullptr: word pointers into shared memory) or on the arenas.
template <bool MASKED>
__device__ __forceinline__ void extend(const SeqArena Q, int64_t qbase, int32_t qrel, int32_t qn, const uint32_t* stQ,
                                       const SeqArena S, int64_t sbase, int32_t srel, int32_t sn, const uint32_t* stS,
                                       int32_t* ring, const ScratchPool& sp, ExtRes* r,
                                       unsigned long long* steps, unsigned long long* retries) {
    if (MASKED) {
        if (!greedy_extend<true>(Q, qbase + qrel, qn, S, sbase + srel, sn, ring, RCQ_RING, r, steps)) {
            if ((threadIdx.x & 31) == 0) (*retries)++;
            const int32_t slot = scratch_acquire(sp);
            greedy_extend<false>(Q, qbase + qrel, qn, S, sbase + srel, sn, sp.rows + slot * sp.per_slot, qn + sn + 3, r, steps);
            scratch_release(sp, slot);
        }
    } else {
        // transcripts start on 32-base boundaries, so word offsets inside a transcript are qrel >> 4
        const int32_t qo = qrel & 15, so = srel & 15;
        if (stQ) {
            SeqShared q, t;
            q.b = stQ + (qrel >> 4); t.b = stS + (srel >> 4);
            if (!greedy_fast<true, SeqShared>(q, qo, qn, t, so, sn, ring, RCQ_RING, r, steps)) {
                if ((threadIdx.x & 31) == 0) (*retries)++;
                greedy_fast_wide<SeqShared>(q, qo, qn, t, so, sn, sp, r, steps);
            }
        } else {
            SeqGlobal q, t;
            q.b = Q.bits + ((qbase + qrel) >> 4); t.b = S.bits + ((sbase + srel) >> 4);
            if (!greedy_fast<true, SeqGlobal>(q, qo, qn, t, so, sn, ring, RCQ_RING, r, steps)) {
                if ((threadIdx.x & 31) == 0) (*retries)++;
                greedy_fast_wide<SeqGlobal>(q, qo, qn, t, so, sn, sp, r, steps);
            }
        }
    }
    __syncwarp();
}

// canonical processing order inside a group: score desc, diag asc, a0 asc, a1 asc
__device__ __forceinline__ bool ung_before(const Ungapped& x, const Ungapped& y) {
    if (x.score != y.score) return x.score > y.score;
    if (x.diag != y.diag) return x.diag < y.diag;
    if (x.a0 != y.a0) return x.a0 < y.a0;
    return x.a1 < y.a1;
}

// Warp per group (dynamic scheduling).  Ungapped HSPs of the group are visited in canonical
// order; one contained in an earlier gapped HSP of the group is skipped; identical results are
// dropped; at most RCQ_MAX_GAPPED_PER_GROUP gapped HSPs are kept.  Group g writes its results
// to out[head .. head+count), the remaining slots of its range keep ta = -1.
// U is the HSP list sorted by group (empty slots, score < 0, sort behind every group);
// keys[] are the sorted group keys of U (k_hsp_keys): ta | tb | strand.
template <bool MASKED>
__global__ void __launch_bounds__(128, 5) k_gapped(const Ungapped* __restrict__ U,
                                                const uint64_t* __restrict__ keys, int64_t n, KeyFmt F,
                                                const uint32_t* __restrict__ heads, int64_t n_groups,
                                                PairView P, Gapped* __restrict__ out, ScratchPool sp,
                                                unsigned long long* __restrict__ ctr /* [0]=next group [1]=cells [2]=retries [3]=n gapped */,
                                                const uint32_t* __restrict__ hlist /* group heads given up by k_gapped_reg, or NULL = all groups */,
                                                const unsigned long long* __restrict__ hcount, unsigned long long* __restrict__ hnext) {
    const int lane = threadIdx.x & 31;
    __shared__ __align__(16) int32_t ring_all[4][6 * RCQ_RING];
    // staged strands of the group's two transcripts (fast path): A forward, A reverse complement, B on the
    // hit strand, B on the opposite strand; RCQ_STAGE_W words each (transcripts up to RCQ_STAGE_MAX bases)
    __shared__ uint32_t stage_all[MASKED ? 1 : 4][MASKED ? 1 : 4 * RCQ_STAGE_W];
    int32_t* ring = ring_all[threadIdx.x >> 5];
    uint32_t* stage = stage_all[MASKED ? 0 : (threadIdx.x >> 5)];
    unsigned long long cells = 0, retries = 0, produced = 0;
    while (true) {
        unsigned long long g = 0;
        if (lane == 0) g = atomicAdd(hlist ? hnext : ctr, 1ull);
        g = __shfl_sync(FULL, g, 0);
        // deferred mode: the list of group heads k_gapped_reg could not finish
        if ((int64_t)g >= (hlist ? (int64_t)*hcount : n_groups)) break;
        const int64_t h = hlist ? hlist[g] : heads[g];
        int64_t e = h + 1;
        const uint64_t key = keys[h];
        while (e < n && keys[e] == key) e++;
        const int32_t gsize = (int32_t)(e - h);
        const int32_t ta = (int32_t)(key >> (F.tb_bits + 1)), tb = (int32_t)((key >> 1) & ((1u << F.tb_bits) - 1u)), minus = (int32_t)(key & 1);
        const int64_t oa = __ldg(P.offA + ta), ob = __ldg(P.offB + tb);
        const int32_t la = __ldg(P.lenA + ta), lb = __ldg(P.lenB + tb);
        const SeqArena Bs = minus ? P.Br : P.Bf, Bo = minus ? P.Bf : P.Br;
        const bool staged = !MASKED && la <= RCQ_STAGE_MAX && lb <= RCQ_STAGE_MAX;
        if (staged) {
            __syncwarp();
            const int32_t wa = (la + 15) / 16 + 2, wb = (lb + 15) / 16 + 2;
            const uint32_t* ga = P.Af.bits + (oa >> 4); const uint32_t* gr = P.Ar.bits + (oa >> 4);
            const uint32_t* gs = Bs.bits + (ob >> 4); const uint32_t* go = Bo.bits + (ob >> 4);
            for (int32_t w = lane; w < wa; w += 32) { stage[w] = __ldg(ga + w); stage[RCQ_STAGE_W + w] = __ldg(gr + w); }
            for (int32_t w = lane; w < wb; w += 32) { stage[2 * RCQ_STAGE_W + w] = __ldg(gs + w); stage[3 * RCQ_STAGE_W + w] = __ldg(go + w); }
            __syncwarp();
        }
        const uint32_t* stAf = staged ? stage : nullptr;
        const uint32_t* stAr = staged ? stage + RCQ_STAGE_W : nullptr;
        const uint32_t* stBs = staged ? stage + 2 * RCQ_STAGE_W : nullptr;
        const uint32_t* stBo = staged ? stage + 3 * RCQ_STAGE_W : nullptr;
        int32_t count = 0;
        Ungapped last; last.score = 0x7fffffff; last.diag = -0x7fffffff; last.a0 = -1; last.a1 = -1;
        bool have_last = false;
        for (int32_t it = 0; it < gsize && count < RCQ_MAX_GAPPED_PER_GROUP; it++) {
            // next ungapped HSP strictly after `last` in canonical order (duplicates collapse)
            Ungapped cur;
            {
                Ungapped bestu; bool found = false;
                for (int64_t x = h + lane; x < e; x += 32) {
                    Ungapped u = U[x];
                    if (u.score < 0) continue;
                    if (have_last && !ung_before(last, u)) continue;
                    if (!found || ung_before(u, bestu)) { bestu = u; found = true; }
                }
                for (int o = 16; o; o >>= 1) {
                    Ungapped v; bool f2;
                    v.score = __shfl_xor_sync(FULL, bestu.score, o); v.diag = __shfl_xor_sync(FULL, bestu.diag, o);
                    v.a0 = __shfl_xor_sync(FULL, bestu.a0, o); v.a1 = __shfl_xor_sync(FULL, bestu.a1, o);
                    v.i0 = __shfl_xor_sync(FULL, bestu.i0, o); f2 = __shfl_xor_sync(FULL, (int)found, o);
                    if (f2 && (!found || ung_before(v, bestu))) { bestu.score = v.score; bestu.diag = v.diag; bestu.a0 = v.a0; bestu.a1 = v.a1; bestu.i0 = v.i0; found = true; }
                }
                if (!found) break;
                cur = bestu; cur.ta = ta; cur.tb = tb; cur.minus = minus;
            }
            last = cur; have_last = true;
            // containment in an earlier gapped HSP of this group
            const int32_t ub0 = cur.a0 + cur.diag, ub1 = cur.a1 + cur.diag;
            bool contained = false;
            if (lane < count) {
                const Gapped gp = out[h + lane];
                contained = gp.a0 <= cur.a0 && cur.a1 <= gp.a1 && gp.b0 <= ub0 && ub1 <= gp.b1;
            }
            if (__any_sync(FULL, contained)) continue;
            const int32_t i0 = cur.i0, j0 = i0 + cur.diag;
            ExtRes L, R;
            extend<MASKED>(P.Ar, oa, la - 1 - i0, i0 + 1, stAr, Bo, ob, lb - 1 - j0, j0 + 1, stBo, ring, sp, &L, &cells, &retries);
            extend<MASKED>(P.Af, oa, i0 + 1, la - i0 - 1, stAf, Bs, ob, j0 + 1, lb - j0 - 1, stBs, ring, sp, &R, &cells, &retries);
            Gapped gp;
            gp.ta = ta; gp.tb = tb; gp.minus = minus;
            gp.a0 = i0 + 1 - L.qi; gp.a1 = i0 + 1 + R.qi;
            gp.b0 = j0 + 1 - L.sj; gp.b1 = j0 + 1 + R.sj;
            gp.score2 = L.score2 + R.score2;
            gp.mism = L.mism + R.mism;
            gp.gaps = (L.gw >> 17) + (R.gw >> 17);
            gp.gapopen = ((L.gw >> 3) & 0x3fff) + ((R.gw >> 3) & 0x3fff);
            gp.nident = ((gp.a1 - gp.a0) + (gp.b1 - gp.b0) - gp.gaps) / 2 - gp.mism;
            bool dup = false;
            if (lane < count) {
                const Gapped o2 = out[h + lane];
                dup = o2.a0 == gp.a0 && o2.a1 == gp.a1 && o2.b0 == gp.b0 && o2.b1 == gp.b1;
            }
            if (__any_sync(FULL, dup)) continue;
            if (lane == 0) out[h + count] = gp;
            __syncwarp();
            count++;
            produced++;
        }
    }
    if (lane == 0) {
        if (cells) atomicAdd(ctr + 1, cells);
        if (retries) atomicAdd(ctr + 2, retries);
        if (produced) atomicAdd(ctr + 3, produced);
    }
}

