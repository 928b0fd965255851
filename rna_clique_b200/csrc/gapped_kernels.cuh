// gapped_kernels.cuh — kernel (3b): X-drop gapped extension with linear gap cost (megablast scoring:
// +1/-2, gap 2.5 per position; kept doubled: +2/-4/-5; X-drop 100 bits = 54 -> 108 doubled).
//
// One warp per (transcript of A, transcript of B, strand) group.  The DP runs in anti-diagonal
// order: lane l owns the two diagonals k = -32+2l and k+1 of a 64-diagonal band, so every
// anti-diagonal step computes one cell per lane from its own registers plus ONE neighbour value
// moved with a warp shuffle (anti-diagonal propagation); the best score so far is shared with
// redux.sync (__reduce_max_sync).  Score and alignment statistics (mismatches, gap characters,
// gap openings, last operation) travel with the cell, so no traceback pass is needed.
// The band is exact as long as no live cell touches its edge; otherwise the extension is redone
// by the unbounded routine below, whose per-diagonal state lives in a global scratch row.
//
// Roofline: integer-ALU bound (no HBM traffic to speak of: sequence words are re-read from L2
// once per 31 cells).  Work unit: DP cells; reported as GCUPS.
#pragma once
#include "search.cuh"

#define NEG (-(1 << 29))
#define FULL 0xffffffffu

struct ExtRes { int32_t score2, qi, sj, mism, gw; };  // gw = gaps<<16 | gapopen<<2 | lastop

__device__ __forceinline__ int32_t gw_gap(int32_t gw, int32_t op) {  // op 1 = U, 2 = L
    return ((gw & ~3) | op) + (1 << 16) + (((gw & 3) != op) ? 4 : 0);
}

__device__ __forceinline__ uint32_t base_at(const uint32_t* bits, int64_t p) {
    return (__ldg(bits + (p >> 4)) >> ((p & 15) * 2)) & 3u;
}
__device__ __forceinline__ uint32_t mask_at(const uint32_t* mask, int64_t p) {
    return mask ? ((__ldg(mask + (p >> 5)) >> (p & 31)) & 1u) : 0u;
}

// ---- fast path: 64-diagonal register band -------------------------------------------------
// q bases at arena positions qpos + i (i in [0, qn)), s bases at spos + j.  Returns false when a
// live cell reached the band edge (result invalid, caller falls back).
template <bool HAS_N>
__device__ bool xdrop_band64(const SeqArena Q, int64_t qpos, int32_t qn, const SeqArena S, int64_t spos,
                             int32_t sn, ExtRes* res, unsigned long long* cells) {
    const int lane = threadIdx.x & 31;
    const int32_t kA = -32 + 2 * lane;
    int32_t hA = (lane == 16) ? 0 : NEG, hB = NEG, mA = 0, mB = 0, gA = 0, gB = 0;
    int32_t bH = (lane == 16) ? 0 : NEG, bI = 0, bJ = 0, bM = 0, bG = 0;
    int32_t gbest = 0;
    uint64_t qw = 0, sw = 0;
    uint32_t qm = 0, sm = 0;
    unsigned long long ncell = 0;
    const int32_t tmax_all = qn + sn;
    bool ok = true;
    for (int32_t p = 0; 2 * p + 1 <= tmax_all; p++) {
        const int32_t c = p - lane + 16;      // iB; iA = c + 1
        const int32_t jB = c + kA + 1;        // = jA
        if (p % 31 == 0) {
            qw = fetch32(Q.bits, qpos + c - 1);
            sw = fetch32(S.bits, spos + jB - 1);
            if (HAS_N) {
                qm = Q.mask ? fetchmask32(Q.mask, qpos + c - 1) : 0u;
                sm = S.mask ? fetchmask32(S.mask, spos + jB - 1) : 0u;
            }
        }
        uint32_t q0 = (uint32_t)qw & 3u, q1 = (uint32_t)(qw >> 2) & 3u, s0 = (uint32_t)sw & 3u;
        bool n0 = false, n1 = false;
        if (HAS_N) { n0 = ((qm | sm) & 1u) != 0; n1 = (((qm >> 1) | sm) & 1u) != 0; }
        qw >>= 2; sw >>= 2;
        if (HAS_N) { qm >>= 1; sm >>= 1; }
        // ---- sub-step 1: t = 2p+1, cells on diagonal kB = kA+1 at (iB, jB) = (c, jB)
        bool any1;
        {
            int32_t hU = __shfl_down_sync(FULL, hA, 1), mU = __shfl_down_sync(FULL, mA, 1),
                    gU = __shfl_down_sync(FULL, gA, 1);
            if (lane == 31) hU = NEG;
            bool valid = c >= 0 && jB >= 0 && c <= qn && jB <= sn;
            int32_t h = NEG, m = 0, g = 0;
            if (valid) {
                ncell++;
                if (c >= 1 && jB >= 1 && hB > NEG) {
                    bool mt = (q0 == s0) && !n0;
                    h = hB + (mt ? RCQ_MATCH2 : RCQ_MISMATCH2); m = mB + (mt ? 0 : 1); g = gB & ~3;
                }
                if (c >= 1 && hU > NEG && hU + RCQ_GAP2 > h) { h = hU + RCQ_GAP2; m = mU; g = gw_gap(gU, 1); }
                if (jB >= 1 && hA > NEG && hA + RCQ_GAP2 > h) { h = hA + RCQ_GAP2; m = mA; g = gw_gap(gA, 2); }
                if (h < gbest - RCQ_XDROP_GAPPED2) h = NEG;
            }
            hB = h; mB = m; gB = g;
            bool alive = h > NEG;
            if (alive && h > bH) { bH = h; bI = c; bJ = jB; bM = m; bG = g; }
            int32_t tm = __reduce_max_sync(FULL, alive ? h : NEG);
            any1 = tm > NEG;
            if (tm > gbest) gbest = tm;
            if (__any_sync(FULL, alive && lane == 31)) ok = false;
        }
        // ---- sub-step 2: t = 2p+2, cells on diagonal kA at (c+1, jB)
        bool any2 = false;
        if (2 * p + 2 <= tmax_all) {
            int32_t hL = __shfl_up_sync(FULL, hB, 1), mL = __shfl_up_sync(FULL, mB, 1),
                    gL = __shfl_up_sync(FULL, gB, 1);
            if (lane == 0) hL = NEG;
            const int32_t iA = c + 1;
            bool valid = iA >= 0 && jB >= 0 && iA <= qn && jB <= sn;
            int32_t h = NEG, m = 0, g = 0;
            if (valid) {
                ncell++;
                if (iA >= 1 && jB >= 1 && hA > NEG) {
                    bool mt = (q1 == s0) && !n1;
                    h = hA + (mt ? RCQ_MATCH2 : RCQ_MISMATCH2); m = mA + (mt ? 0 : 1); g = gA & ~3;
                }
                if (iA >= 1 && hB > NEG && hB + RCQ_GAP2 > h) { h = hB + RCQ_GAP2; m = mB; g = gw_gap(gB, 1); }
                if (jB >= 1 && hL > NEG && hL + RCQ_GAP2 > h) { h = hL + RCQ_GAP2; m = mL; g = gw_gap(gL, 2); }
                if (h < gbest - RCQ_XDROP_GAPPED2) h = NEG;
            }
            hA = h; mA = m; gA = g;
            bool alive = h > NEG;
            if (alive && h > bH) { bH = h; bI = iA; bJ = jB; bM = m; bG = g; }
            int32_t tm = __reduce_max_sync(FULL, alive ? h : NEG);
            any2 = tm > NEG;
            if (tm > gbest) gbest = tm;
            if (__any_sync(FULL, alive && lane == 0)) ok = false;
        }
        if (!ok) break;
        if (!any1 && !any2) break;
    }
    for (int o = 16; o; o >>= 1) ncell += __shfl_down_sync(FULL, ncell, o);
    if (lane == 0) *cells += ncell;
    if (!ok) return false;
    // best end point: score desc, then t asc, then k asc
    uint64_t key = 0;
    if (bH > NEG) {
        int32_t t = bI + bJ, k = bJ - bI;
        key = ((uint64_t)(uint32_t)(bH + (1 << 20)) << 42) | ((uint64_t)((1 << 22) - 1 - t) << 20) |
              (uint64_t)((1 << 20) - 1 - (k + (1 << 19)));
    }
    uint64_t best = key;
    for (int o = 16; o; o >>= 1) { uint64_t x = __shfl_xor_sync(FULL, best, o); if (x > best) best = x; }
    int src = __ffs(__ballot_sync(FULL, key == best)) - 1;
    res->score2 = __shfl_sync(FULL, bH, src);
    res->qi = __shfl_sync(FULL, bI, src);
    res->sj = __shfl_sync(FULL, bJ, src);
    res->mism = __shfl_sync(FULL, bM, src);
    res->gw = __shfl_sync(FULL, bG, src);
    return true;
}

// ---- general path: unbounded live window, per-diagonal state in a global scratch row -------
// scratch: 3 * (qn + sn + 3) int32.  Same recurrence, same tie rules, same pruning.
__device__ void xdrop_wide(const SeqArena Q, int64_t qpos, int32_t qn, const SeqArena S, int64_t spos,
                           int32_t sn, int32_t* scratch, ExtRes* res, unsigned long long* cells) {
    const int lane = threadIdx.x & 31;
    const int32_t nd = qn + sn + 3;
    int32_t *H = scratch, *MM = scratch + nd, *GW = scratch + 2 * nd;
    for (int32_t x = lane; x < nd; x += 32) { H[x] = NEG; MM[x] = 0; GW[x] = 0; }
    __syncwarp();
#define IX(k) ((k) + qn + 1)
    if (lane == 0) H[IX(0)] = 0;
    __syncwarp();
    int32_t bH = (lane == 0) ? 0 : NEG, bI = 0, bJ = 0, bM = 0, bG = 0;
    int32_t gbest = 0, lo1 = 0, hi1 = 0, lo2 = 1, hi2 = 0;
    unsigned long long ncell = 0;
    for (int32_t t = 1; t <= qn + sn; t++) {
        if (lo1 > hi1 && lo2 > hi2) break;
        int32_t lo, hi;
        if (lo1 > hi1) { lo = lo2; hi = hi2; }
        else if (lo2 > hi2) { lo = lo1; hi = hi1; }
        else { lo = min(lo1, lo2); hi = max(hi1, hi2); }
        lo = max(lo - 1, -qn); hi = min(hi + 1, sn);
        int32_t nlo = 1 << 28, nhi = -(1 << 28), tm = NEG;
        int32_t k0 = lo + ((t - lo) & 1);
        for (int32_t k = k0 + 2 * lane; k <= hi; k += 64) {
            int32_t i = (t - k) >> 1, j = (t + k) >> 1;
            if (i < 0 || j < 0 || i > qn || j > sn) continue;
            ncell++;
            int32_t h = NEG, m = 0, g = 0;
            int32_t hd = H[IX(k)], hu = H[IX(k + 1)], hl = H[IX(k - 1)];
            if (i >= 1 && j >= 1 && hd > NEG) {
                bool mt = base_at(Q.bits, qpos + i - 1) == base_at(S.bits, spos + j - 1) &&
                          !mask_at(Q.mask, qpos + i - 1) && !mask_at(S.mask, spos + j - 1);
                h = hd + (mt ? RCQ_MATCH2 : RCQ_MISMATCH2); m = MM[IX(k)] + (mt ? 0 : 1); g = GW[IX(k)] & ~3;
            }
            if (i >= 1 && hu > NEG && hu + RCQ_GAP2 > h) { h = hu + RCQ_GAP2; m = MM[IX(k + 1)]; g = gw_gap(GW[IX(k + 1)], 1); }
            if (j >= 1 && hl > NEG && hl + RCQ_GAP2 > h) { h = hl + RCQ_GAP2; m = MM[IX(k - 1)]; g = gw_gap(GW[IX(k - 1)], 2); }
            if (h > NEG && h < gbest - RCQ_XDROP_GAPPED2) h = NEG;
            H[IX(k)] = h; MM[IX(k)] = m; GW[IX(k)] = g;
            if (h > NEG) {
                nlo = min(nlo, k); nhi = max(nhi, k); tm = max(tm, h);
                if (h > bH) { bH = h; bI = i; bJ = j; bM = m; bG = g; }
            }
        }
        __syncwarp();
        tm = __reduce_max_sync(FULL, tm);
        nlo = __reduce_min_sync(FULL, nlo);
        nhi = __reduce_max_sync(FULL, nhi);
        if (tm > gbest) gbest = tm;
        lo2 = lo1; hi2 = hi1;
        if (tm > NEG) { lo1 = nlo; hi1 = nhi; } else { lo1 = 1; hi1 = 0; }
    }
#undef IX
    for (int o = 16; o; o >>= 1) ncell += __shfl_down_sync(FULL, ncell, o);
    if (lane == 0) *cells += ncell;
    uint64_t key = 0;
    if (bH > NEG) {
        int32_t t = bI + bJ, k = bJ - bI;
        key = ((uint64_t)(uint32_t)(bH + (1 << 20)) << 42) | ((uint64_t)((1 << 22) - 1 - t) << 20) |
              (uint64_t)((1 << 20) - 1 - (k + (1 << 19)));
    }
    uint64_t best = key;
    for (int o = 16; o; o >>= 1) { uint64_t x = __shfl_xor_sync(FULL, best, o); if (x > best) best = x; }
    int src = __ffs(__ballot_sync(FULL, key == best)) - 1;
    res->score2 = __shfl_sync(FULL, bH, src);
    res->qi = __shfl_sync(FULL, bI, src);
    res->sj = __shfl_sync(FULL, bJ, src);
    res->mism = __shfl_sync(FULL, bM, src);
    res->gw = __shfl_sync(FULL, bG, src);
}

// ---- one extension, fast path with fallback -------------------------------------------------
template <bool HAS_N>
__device__ __forceinline__ void extend(const SeqArena Q, int64_t qpos, int32_t qn, const SeqArena S,
                                       int64_t spos, int32_t sn, int32_t* scratch, ExtRes* r,
                                       unsigned long long* cells, unsigned long long* retries) {
    if (!xdrop_band64<HAS_N>(Q, qpos, qn, S, spos, sn, r, cells)) {
        if ((threadIdx.x & 31) == 0) (*retries)++;
        xdrop_wide(Q, qpos, qn, S, spos, sn, scratch, r, cells);
    }
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
template <bool HAS_N>
__global__ void __launch_bounds__(128) k_gapped(const Ungapped* __restrict__ U, const uint32_t* __restrict__ order,
                                                const uint64_t* __restrict__ keys, int64_t n,
                                                const uint32_t* __restrict__ heads, int64_t n_groups,
                                                PairView P, Gapped* __restrict__ out,
                                                int32_t* __restrict__ scratch, int64_t scratch_per_warp,
                                                unsigned long long* __restrict__ ctr /* [0]=next group [1]=cells [2]=retries [3]=n gapped */) {
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int32_t* my_scratch = scratch + (int64_t)warp_global * scratch_per_warp;
    unsigned long long cells = 0, retries = 0, produced = 0;
    while (true) {
        unsigned long long g = 0;
        if (lane == 0) g = atomicAdd(ctr, 1ull);
        g = __shfl_sync(FULL, g, 0);
        if ((int64_t)g >= n_groups) break;
        const int64_t h = heads[g];
        int64_t e = h + 1;
        const uint64_t key = keys[h];
        while (e < n && keys[e] == key) e++;
        const int32_t gsize = (int32_t)(e - h);
        const Ungapped first = U[order[h]];
        const int32_t ta = first.ta, tb = first.tb, minus = first.minus;
        const int64_t oa = __ldg(P.offA + ta), ob = __ldg(P.offB + tb);
        const int32_t la = __ldg(P.lenA + ta), lb = __ldg(P.lenB + tb);
        const SeqArena Bs = minus ? P.Br : P.Bf, Bo = minus ? P.Bf : P.Br;
        int32_t count = 0;
        Ungapped last; last.score = 0x7fffffff; last.diag = -0x7fffffff; last.a0 = -1; last.a1 = -1;
        bool have_last = false;
        for (int32_t it = 0; it < gsize && count < RCQ_MAX_GAPPED_PER_GROUP; it++) {
            // next ungapped HSP strictly after `last` in canonical order (duplicates collapse)
            Ungapped cur;
            if (gsize == 1) {
                cur = first;
            } else {
                Ungapped bestu; bool found = false;
                for (int64_t x = h + lane; x < e; x += 32) {
                    Ungapped u = U[order[x]];
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
            extend<HAS_N>(P.Ar, oa + (la - 1 - i0), i0 + 1, Bo, ob + (lb - 1 - j0), j0 + 1, my_scratch, &L, &cells, &retries);
            extend<HAS_N>(P.Af, oa + i0 + 1, la - i0 - 1, Bs, ob + j0 + 1, lb - j0 - 1, my_scratch, &R, &cells, &retries);
            Gapped gp;
            gp.ta = ta; gp.tb = tb; gp.minus = minus;
            gp.a0 = i0 + 1 - L.qi; gp.a1 = i0 + 1 + R.qi;
            gp.b0 = j0 + 1 - L.sj; gp.b1 = j0 + 1 + R.sj;
            gp.score2 = L.score2 + R.score2;
            gp.mism = L.mism + R.mism;
            gp.gaps = (L.gw >> 16) + (R.gw >> 16);
            gp.gapopen = ((L.gw >> 2) & 0x3fff) + ((R.gw >> 2) & 0x3fff);
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

