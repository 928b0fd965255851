// gapped_half.cuh — kernel (3b), main path: the greedy X-drop gapped extension of gapped_kernels.cuh (same
// recurrence, same results) organised around HALF-WARPS with the wave rows in registers.
//
// Why: the extension is bound by instruction issue, and a wave of the recurrence is narrow - 1 diagonal at the
// start, about 2 x 19 + 1 once BLAST's X-drop window has opened.  One diagonal per lane and rows in shared
// memory (k_gapped) spent two 32-lane passes per wave with 7 lanes busy in the second, and most instructions on
// addressing, range checks and sentinels.  Here
//   * a half-warp (16 lanes) owns one (transcript, transcript, strand) group and is an independent extension
//     engine; the two halves of a warp run different groups through ONE instruction stream;
//   * every lane owns three ADJACENT diagonals, 48 per half-warp, held in registers: furthest row R, mismatch
//     count M and gap word G of the previous wave.  The two neighbours a lane lacks come from two shuffles, so a
//     wave needs no shared-memory rows, no stores, no sentinels and no index arithmetic;
//   * the three candidate entries carry their tie rank in the two low bits (row << 2 | rank), so the entry and
//     the operation that made it come out of one three-way maximum (VIMNMX3); candidates beyond a sequence end
//     exceed the row limit by exactly one, so one compare sends the (rare) end-of-sequence case to the
//     full per-candidate clamps;
//   * the 48-diagonal window follows the band: it is re-centred by whole lanes when the band drifts (indels); a
//     band wider than 48, or transcripts longer than the staged limit, defer the WHOLE group to k_gapped (global
//     rows of any width), which then runs on the list of deferred groups.
// The main loop is a state machine: the divergent part (fetch a group, stage its four strands in shared memory,
// pick the next ungapped HSP in canonical order, containment / duplicate checks, reduce the best end point) runs
// per half-warp at the top of the loop; the wave step below it is executed by both halves together.
//
// Roofline: instruction issue (INT32); HBM traffic is the transcripts once per group.  Work unit: extension
// steps (base comparisons), counted exactly as the oracle counts them.
#pragma once
#include "gapped_kernels.cuh"

#define GH_THREADS 128
#ifndef GH_MIN_BLOCKS
#define GH_MIN_BLOCKS 4               // resident blocks per SM the register budget is set for
#endif

enum { GH_FETCH = 0, GH_NEXT = 1, GH_RUN = 2, GH_FIN = 3, GH_DONE = 4 };

// both halves of the warp slide along one diagonal each: number of matching base pairs from (xq, xs), at most
// lim (lim <= 0: nothing to do for this half).  16 lanes x 16 bases per round.
__device__ __forceinline__ int32_t gh_coop_slide(const uint32_t* Q, int32_t xq, const uint32_t* S, int32_t xs, int32_t lim,
                                                 int j, int half) {
    int32_t base = 0, result = lim > 0 ? lim : 0;
    bool done = lim <= 0;
    while (true) {
        const int32_t o = base + 16 * j;
        int32_t z = 0;
        if (!done && o < lim) {
            SeqShared q, s;
            q.b = Q; s.b = S;
            z = min(run16(q, xq + o, s, xs + o), lim - o);
        }
        const unsigned stop = (__ballot_sync(FULL, z < 16) >> (16 * half)) & 0xFFFFu;
        const int src = stop ? __ffs(stop) - 1 : 0;
        const int32_t zs = __shfl_sync(FULL, z, src, 16);
        if (!done && stop) { result = base + 16 * src + zs; done = true; }
        base += 256;
        if (!done && base >= lim) { result = lim; done = true; }
        if (__all_sync(FULL, done)) break;
    }
    return result;
}

struct GhExt {            // one extension of a half-warp (uniform over its 16 lanes)
    const uint32_t *Q, *S;     // staged words of the two strands, already advanced to the word of the start base
    int32_t qo, so;            // offset of the start base inside that word (0 .. 15)
    int32_t qn, sn;
};

// CELLS adjacent diagonals per lane: 3 (48-diagonal window) for the first pass over all groups, 4 (64) for the
// groups the first pass gave up on; what outgrows that too goes to k_gapped.  Groups come from `heads` (in_list
// == NULL) or from a list of group heads written by an earlier pass; groups given up are appended to out_list.
template <int CELLS>
__global__ void __launch_bounds__(GH_THREADS, GH_MIN_BLOCKS) k_gapped_half(
    const Ungapped* __restrict__ U, const uint64_t* __restrict__ keys, int64_t n, KeyFmt F,
    const uint32_t* __restrict__ heads, int64_t n_groups, PairView P, Gapped* __restrict__ out,
    const uint32_t* __restrict__ in_list, const unsigned long long* __restrict__ in_count, unsigned long long* __restrict__ in_next,
    uint32_t* __restrict__ out_list, unsigned long long* __restrict__ out_count,
    unsigned long long* __restrict__ ctr /* [1]=cells [2]=groups given up by the first pass [3]=n gapped */) {
    constexpr int GH_CELLS = CELLS;
    constexpr int GH_WIN = 16 * CELLS;           // diagonals per half-warp window
    constexpr int GH_KB0 = -(GH_WIN / 2);        // window start of a new extension: diagonal 0 in the middle
    __shared__ uint32_t stage_all[GH_THREADS / 16][4 * RCQ_STAGE_W];
    const int lane = threadIdx.x & 31, half = lane >> 4, j = lane & 15;
    const int hs = 16 * half;                       // shift of this half's bits in a ballot
    uint32_t* stage = stage_all[threadIdx.x >> 4];
    const uint32_t* stAf = stage;
    const uint32_t* stAr = stage + RCQ_STAGE_W;
    const uint32_t* stBs = stage + 2 * RCQ_STAGE_W;
    const uint32_t* stBo = stage + 3 * RCQ_STAGE_W;

    // Every warp-wide primitive below runs with the FULL mask at a point all 32 lanes reach together (branches
    // around them test ballot results, which are uniform): the two halves must never split into two sub-warps,
    // or each would pay for the whole instruction stream.  A half that has nothing to do in a block takes part
    // with its predicate off.
    int st = GH_FETCH;
    // group
    int64_t h = 0, e_ = 0;
    int32_t gsize = 0, ta = 0, tb = 0, minus = 0, la = 0, lb = 0, count = 0, it = 0;
    Ungapped last; last.score = 0; last.diag = 0; last.a0 = 0; last.a1 = 0; last.i0 = 0; last.ta = last.tb = last.minus = 0;
    bool have_last = false, group_open = false;
    // current HSP / extension
    int32_t i0 = 0, j0 = 0, phase = 0;
    GhExt X; X.Q = stage; X.S = stage; X.qo = X.so = 0; X.qn = X.sn = 0;
    int32_t Lscore2 = 0, Lqi = 0, Lsj = 0, Lmism = 0, Lgw = 0;
    // wave state
    int32_t d = 0, Lk = 0, Uk = 0, gb = 0, Tprev = 0, kb = GH_KB0;
    int32_t R[CELLS];
    uint32_t M[CELLS], G[CELLS];
#pragma unroll
    for (int c = 0; c < CELLS; c++) { R[c] = NEG; M[c] = 0; G[c] = 0; }
    int32_t bS = NEG, bD = 0, bK = 0, bM = 0, bG = 0;
    int32_t T0 = 0, T1 = 0;                 // T(x) for x & 31 == j and == j + 16
    bool overflow = false;
    uint32_t nstep = 0;                     // extension steps of the open group (this lane's share)
    unsigned long long cells = 0, produced = 0, n_deferred = 0;

    while (true) {
        // ------------------------------------------------------------------ set-up passes (both halves together)
        while (__any_sync(FULL, st != GH_RUN && st != GH_DONE)) {
            bool begin = false;                       // this half starts an extension at the end of the pass
            const uint32_t* bQ = stage; const uint32_t* bS_ = stage;
            int32_t bqrel = 0, bqn = 0, bsrel = 0, bsn = 0;
            // ---- (1) an extension has ended
            if (__any_sync(FULL, st == GH_FIN)) {
                bool p = st == GH_FIN;
                if (p && overflow) {
                    // the band outgrew the window: the whole group goes to the wide kernel
                    if (j == 0) out_list[atomicAdd(out_count, 1ull)] = (uint32_t)h;
                    n_deferred++;
                    group_open = false; nstep = 0; overflow = false;
                    st = GH_FETCH; p = false;
                }
                // best end point of the extension: score desc, then d asc, then k asc
                uint64_t key = 0;
                if (p && bS > NEG)
                    key = ((uint64_t)(uint32_t)(bS + (1 << 20)) << 42) | ((uint64_t)((1 << 21) - 1 - bD) << 21) |
                          (uint64_t)((1 << 21) - 1 - (bK + (1 << 20)));
                uint64_t best = key;
#pragma unroll
                for (int o = 8; o; o >>= 1) { const uint64_t x = __shfl_xor_sync(FULL, best, o, 16); if (x > best) best = x; }
                const unsigned who = (__ballot_sync(FULL, key == best) >> hs) & 0xFFFFu;
                const int src = who ? __ffs(who) - 1 : 0;
                const int32_t rS = __shfl_sync(FULL, bS, src, 16), rD = __shfl_sync(FULL, bD, src, 16),
                              rK = __shfl_sync(FULL, bK, src, 16), rM = __shfl_sync(FULL, bM, src, 16),
                              rG = __shfl_sync(FULL, bG, src, 16);
                const int32_t rI = (rS - rK + 6 * rD) >> 1;          // score = 2 row + k - 6 d
                bool dup = false;
                Gapped gp;
                gp.ta = ta; gp.tb = tb; gp.minus = minus;
                gp.a0 = i0 + 1 - Lqi; gp.a1 = i0 + 1 + rI;
                gp.b0 = j0 + 1 - Lsj; gp.b1 = j0 + 1 + rI + rK;
                gp.score2 = Lscore2 + rS;
                gp.mism = Lmism + rM;
                gp.gaps = (Lgw >> 17) + (rG >> 17);
                gp.gapopen = ((Lgw >> 3) & 0x3fff) + ((rG >> 3) & 0x3fff);
                gp.nident = ((gp.a1 - gp.a0) + (gp.b1 - gp.b0) - gp.gaps) / 2 - gp.mism;
                const bool fin_r = p && phase == 1;
                if (fin_r) {
                    for (int32_t c0 = j; c0 < count; c0 += 16) {
                        const Gapped o2 = out[h + c0];
                        dup = dup || (o2.a0 == gp.a0 && o2.a1 == gp.a1 && o2.b0 == gp.b0 && o2.b1 == gp.b1);
                    }
                }
                const bool any_dup = (__ballot_sync(FULL, dup) >> hs) & 0xFFFFu;
                if (fin_r) {
                    if (!any_dup) {
                        if (j == 0) out[h + count] = gp;
                        count++;
                    }
                    st = GH_NEXT;
                } else if (p) {           // the left part is done: keep it, start the right part
                    Lscore2 = rS; Lqi = rI; Lsj = rI + rK; Lmism = rM; Lgw = rG;
                    phase = 1;
                    begin = true;
                    bQ = stAf; bqrel = i0 + 1; bqn = la - i0 - 1; bS_ = stBs; bsrel = j0 + 1; bsn = lb - j0 - 1;
                }
                __syncwarp();             // out[] written by lane 0 of a half is read by its other lanes below
            }
            // ---- (2) take the next group and stage its four strands
            if (__any_sync(FULL, st == GH_FETCH)) {
                bool p = st == GH_FETCH;
                if (p && group_open) {        // the finished group's work counts; a deferred one was dropped above
                    cells += nstep;
                    produced += (j == 0) ? (unsigned long long)count : 0ull;
                    nstep = 0;
                    group_open = false;
                }
                unsigned long long g = 0;
                if (p && j == 0) g = atomicAdd(in_next, 1ull);
                g = __shfl_sync(FULL, g, 0, 16);
                if (p && (int64_t)g >= (in_list ? (int64_t)*in_count : n_groups)) { st = GH_DONE; p = false; }
                if (p) {
                    h = in_list ? in_list[g] : heads[g];
                    e_ = h + 1;
                    const uint64_t key = keys[h];
                    while (e_ < n && keys[e_] == key) e_++;
                    gsize = (int32_t)(e_ - h);
                    ta = (int32_t)(key >> (F.tb_bits + 1));
                    tb = (int32_t)((key >> 1) & ((1u << F.tb_bits) - 1u));
                    minus = (int32_t)(key & 1);
                    la = __ldg(P.lenA + ta); lb = __ldg(P.lenB + tb);
                    if (la > RCQ_STAGE_MAX || lb > RCQ_STAGE_MAX) {
                        if (j == 0) out_list[atomicAdd(out_count, 1ull)] = (uint32_t)h;
                        n_deferred++;
                        p = false;            // stays in GH_FETCH: next pass takes another group
                    }
                }
                if (p) {
                    const int64_t oa = __ldg(P.offA + ta), ob = __ldg(P.offB + tb);
                    const SeqArena Bs = minus ? P.Br : P.Bf, Bo = minus ? P.Bf : P.Br;
                    const int32_t wa = (la + 15) / 16 + 2, wb = (lb + 15) / 16 + 2;
                    const uint32_t* ga = P.Af.bits + (oa >> 4); const uint32_t* gr = P.Ar.bits + (oa >> 4);
                    const uint32_t* gs = Bs.bits + (ob >> 4); const uint32_t* go = Bo.bits + (ob >> 4);
                    for (int32_t w = j; w < wa; w += 16) { stage[w] = __ldg(ga + w); stage[RCQ_STAGE_W + w] = __ldg(gr + w); }
                    for (int32_t w = j; w < wb; w += 16) { stage[2 * RCQ_STAGE_W + w] = __ldg(gs + w); stage[3 * RCQ_STAGE_W + w] = __ldg(go + w); }
                    count = 0; it = 0; have_last = false; group_open = true;
                    st = GH_NEXT;
                }
                __syncwarp();             // the staged words are read by the other lanes of the half
            }
            // ---- (3) the next ungapped HSP of the group, strictly after `last` in canonical order
            if (__any_sync(FULL, st == GH_NEXT)) {
                bool p = st == GH_NEXT;
                if (p && (it >= gsize || count >= RCQ_MAX_GAPPED_PER_GROUP)) { st = GH_FETCH; p = false; }
                if (p) it++;
                Ungapped bestu; bool found = false;
                bestu.score = 0; bestu.diag = 0; bestu.a0 = 0; bestu.a1 = 0; bestu.i0 = 0;
                if (p) {
                    for (int64_t x = h + j; x < e_; x += 16) {
                        const Ungapped u = U[x];
                        if (u.score < 0) continue;
                        if (have_last && !ung_before(last, u)) continue;
                        if (!found || ung_before(u, bestu)) { bestu = u; found = true; }
                    }
                }
#pragma unroll
                for (int o = 8; o; o >>= 1) {
                    Ungapped v;
                    v.score = __shfl_xor_sync(FULL, bestu.score, o, 16); v.diag = __shfl_xor_sync(FULL, bestu.diag, o, 16);
                    v.a0 = __shfl_xor_sync(FULL, bestu.a0, o, 16); v.a1 = __shfl_xor_sync(FULL, bestu.a1, o, 16);
                    v.i0 = __shfl_xor_sync(FULL, bestu.i0, o, 16);
                    const bool f2 = __shfl_xor_sync(FULL, (int)found, o, 16);
                    if (f2 && (!found || ung_before(v, bestu))) {
                        bestu.score = v.score; bestu.diag = v.diag; bestu.a0 = v.a0; bestu.a1 = v.a1; bestu.i0 = v.i0; found = true;
                    }
                }
                if (p && !found) { st = GH_FETCH; p = false; }
                bool contained = false;
                if (p) {
                    last = bestu; have_last = true;
                    const int32_t ub0 = bestu.a0 + bestu.diag, ub1 = bestu.a1 + bestu.diag;
                    for (int32_t c0 = j; c0 < count; c0 += 16) {
                        const Gapped gq = out[h + c0];
                        contained = contained || (gq.a0 <= bestu.a0 && bestu.a1 <= gq.a1 && gq.b0 <= ub0 && ub1 <= gq.b1);
                    }
                }
                const bool any_contained = (__ballot_sync(FULL, contained) >> hs) & 0xFFFFu;
                if (p && !any_contained) {
                    i0 = bestu.i0; j0 = i0 + bestu.diag;
                    phase = 0;
                    // left part: the anchor pair first, walking backwards = forwards on the opposite strands
                    begin = true;
                    bQ = stAr; bqrel = la - 1 - i0; bqn = i0 + 1; bS_ = stBo; bsrel = lb - 1 - j0; bsn = j0 + 1;
                }
            }
            // ---- (4) start the extensions asked for above: wave 0 and the register rows
            if (__any_sync(FULL, begin)) {
                if (begin) {
                    X.Q = bQ + (bqrel >> 4); X.qo = bqrel & 15; X.qn = bqn;
                    X.S = bS_ + (bsrel >> 4); X.so = bsrel & 15; X.sn = bsn;
                }
                const int32_t first = gh_coop_slide(X.Q, X.qo, X.S, X.so, begin ? min(X.qn, X.sn) : 0, j, half);
                if (begin) {
                    if (j == 0) nstep += (uint32_t)(first + 1);
                    kb = GH_KB0;
                    const bool mine = (0 - GH_KB0) / GH_CELLS == j;     // diagonal 0 = cell (0 - kb) % CELLS of lane (0 - kb) / CELLS
#pragma unroll
                    for (int c = 0; c < CELLS; c++) { R[c] = (mine && c == (0 - GH_KB0) % GH_CELLS) ? first : NEG; M[c] = 0; G[c] = 0; }
                    bS = mine ? 2 * first : NEG; bD = 0; bK = 0; bM = 0; bG = 0;
                    gb = 2 * first; Tprev = gb;
                    T0 = (j == 0) ? gb : 0; T1 = 0;
                    Lk = 0; Uk = 0; d = 0;
                    overflow = false;
                    st = (X.qn + X.sn >= 1) ? GH_RUN : GH_FIN;
                }
            }
        }
        if (__all_sync(FULL, st == GH_DONE)) break;

        // ---------------------------------------------------------------------- one wave, both halves
        const bool act0 = st == GH_RUN;
        d += act0 ? 1 : 0;
        const int32_t qn = X.qn, sn = X.sn;
        int32_t tref;
        {
            const int32_t slot = (d - RCQ_XDROP_OFFSET) & 31;
            // lane (slot & 15) of each half holds the entry; the register is chosen on the receiving side
            const int32_t a0v = __shfl_sync(FULL, T0, slot & 15, 16), a1v = __shfl_sync(FULL, T1, slot & 15, 16);
            tref = (slot & 16) ? a1v : a0v;
            if (d < RCQ_XDROP_OFFSET) tref = 0;
        }
        const int32_t klo = max(Lk - 1, -qn), khi = min(Uk + 1, sn);
        bool act = act0;
        // keep the band inside the window: re-centre by whole lanes, or give the group up when it is wider than the window
        {
            const bool out_of_window = act && (klo < kb || khi > kb + GH_WIN - 1);
            if (__any_sync(FULL, out_of_window)) {
                int32_t sh = 0;                  // lanes to shift: new lane j takes the cells of old lane j + sh
                if (out_of_window) {
                    const int32_t width = khi - klo + 1;
                    if (width > GH_WIN - 2 * GH_CELLS + 2) {       // no room to re-centre with whole-lane shifts
                        overflow = true; st = GH_FIN; act = false;
                    } else {
                        const int32_t want = klo - (GH_WIN - width) / 2;     // ideal window start
                        const int32_t diff = want - kb;
                        sh = diff >= 0 ? diff / GH_CELLS : -((-diff + GH_CELLS - 1) / GH_CELLS);
                        kb += GH_CELLS * sh;
                    }
                }
                const int srcl = j + sh;
                const bool in = srcl >= 0 && srcl < 16;
                const int sl = in ? srcl : j;
#pragma unroll
                for (int c = 0; c < CELLS; c++) {
                    const int32_t r = __shfl_sync(FULL, R[c], sl, 16);
                    M[c] = __shfl_sync(FULL, M[c], sl, 16);
                    G[c] = __shfl_sync(FULL, G[c], sl, 16);
                    R[c] = in ? r : NEG;
                }
            }
        }
        const int32_t d6 = 6 * d;
        // alive iff 2 * row + k >= xcut (dead predecessors are far below any cut); a half that is not running gets a
        // cut nothing passes
        const int32_t xcut = act ? tref - RCQ_XDROP_GAPPED2 + d6 : 0x7fffffff;
        // predecessor rows / counts of cells -1 .. CELLS: the two outer ones come from the neighbour lanes
        int32_t Rx[CELLS + 2];
        uint32_t Mx[CELLS + 2], Gx[CELLS + 2];
        Rx[0] = __shfl_up_sync(FULL, R[CELLS - 1], 1, 16); Rx[CELLS + 1] = __shfl_down_sync(FULL, R[0], 1, 16);
        Mx[0] = __shfl_up_sync(FULL, M[CELLS - 1], 1, 16); Mx[CELLS + 1] = __shfl_down_sync(FULL, M[0], 1, 16);
        Gx[0] = __shfl_up_sync(FULL, G[CELLS - 1], 1, 16); Gx[CELLS + 1] = __shfl_down_sync(FULL, G[0], 1, 16);
        if (j == 0) Rx[0] = NEG;
        if (j == 15) Rx[CELLS + 1] = NEG;
#pragma unroll
        for (int c = 0; c < CELLS; c++) { Rx[c + 1] = R[c]; Mx[c + 1] = M[c]; Gx[c + 1] = G[c]; }
        const int32_t k0 = kb + GH_CELLS * j;
        SeqShared Qs, Ss;
        Qs.b = X.Q; Ss.b = X.S;

        int32_t lim_[CELLS], zsum = 0;
        bool cont[CELLS], anycont = false;
#pragma unroll
        for (int c = 0; c < CELLS; c++) {
            const int32_t k = k0 + c;
            const int32_t left = Rx[c], mid = Rx[c + 1], right = Rx[c + 2];
            const int32_t rowmax = min(qn, sn - k);
            // candidates with their tie rank in the low bits: mismatch (rank 2), query gap (1), subject gap (0)
            const int32_t m4 = __vimax3_s32(mid * 4 + 6, right * 4 + 5, left * 4);
            int32_t bi = m4 >> 2, t = m4 & 3;
            if (bi > rowmax) {
                // an end of a sequence is in reach: candidates past it are invalid one by one
                int32_t c0 = mid + 1, c1 = right + 1, c2 = left;
                if (c0 > rowmax) c0 = NEG;
                if (c1 > rowmax) c1 = NEG;
                if (c2 > rowmax) c2 = NEG;
                bi = c0; t = 2;
                if (c1 > bi) { bi = c1; t = 1; }
                if (c2 > bi) { bi = c2; t = 0; }
            }
            const bool ok = 2 * bi + k >= xcut;
            const bool mis = t == 2, qgap = t == 1;
            const uint32_t sm = mis ? Mx[c + 1] : (qgap ? Mx[c + 2] : Mx[c]);
            const uint32_t sg = mis ? Gx[c + 1] : (qgap ? Gx[c + 2] : Gx[c]);
            const uint32_t op = 2u - (uint32_t)t;                 // 0 mismatch, 1 query gap, 2 subject gap
            const uint32_t base = sg & ~7u;
            const uint32_t gapw = base + op + (1u << 17) + (((sg & 7u) != op) ? 8u : 0u);
            const int32_t lim = rowmax - bi;
            // dead cells compare in-bounds words of the strands and drop the result
            const int32_t bs = max(bi, 0);
            const int32_t r16 = run16(Qs, X.qo + bs, Ss, max(X.so + bs + k, 0));
            const int32_t z = ok ? min(r16, max(lim, 0)) : 0;
            zsum += z;
            lim_[c] = lim;
            R[c] = ok ? bi + z : NEG;
            M[c] = sm + (mis ? 1u : 0u);
            G[c] = (mis ? base : gapw) | (z > 0 ? 4u : 0u);
            cont[c] = z == 16 && lim > 16;
            anycont = anycont || cont[c];
        }
        // slides that survived their first 16 bases are finished by the half-warp together
        if (__any_sync(FULL, anycont)) {
#pragma unroll
            for (int c = 0; c < CELLS; c++) {
                unsigned more = __ballot_sync(FULL, cont[c]);
                while (more) {
                    const unsigned mh = (more >> hs) & 0xFFFFu;
                    const int sl = mh ? __ffs(mh) - 1 : 0;
                    const int32_t e2 = __shfl_sync(FULL, R[c], sl, 16), l2 = __shfl_sync(FULL, lim_[c], sl, 16);
                    const int32_t k2 = kb + GH_CELLS * sl + c;
                    const int32_t extra = gh_coop_slide(X.Q, X.qo + e2, X.S, X.so + e2 + k2, mh ? l2 - 16 : 0, j, half);
                    if (mh && j == sl) { R[c] += extra; zsum += extra; cont[c] = false; }
                    more = __ballot_sync(FULL, cont[c]);
                }
            }
        }
        int32_t wbest = NEG, nL = 1 << 28, nU = -(1 << 28), nalive = 0;
#pragma unroll
        for (int c = 0; c < CELLS; c++) {
            const int32_t k = k0 + c;
            const bool ok = R[c] >= 0;
            const int32_t sc = ok ? 2 * R[c] + k - d6 : NEG;
            wbest = max(wbest, sc);
            nL = min(nL, ok ? k : (1 << 28)); nU = max(nU, ok ? k : -(1 << 28));
            nalive += ok ? 1 : 0;
            if (sc > bS) { bS = sc; bD = d; bK = k; bM = (int32_t)M[c]; bG = (int32_t)G[c]; }
        }
        nstep += (uint32_t)(zsum + nalive);       // one step per base slid over plus one per live cell
#pragma unroll
        for (int o = 8; o; o >>= 1) {
            wbest = max(wbest, __shfl_xor_sync(FULL, wbest, o, 16));
            nL = min(nL, __shfl_xor_sync(FULL, nL, o, 16));
            nU = max(nU, __shfl_xor_sync(FULL, nU, o, 16));
        }
        if (act) {
            Tprev = max(Tprev, wbest);
            if (j == (d & 15)) { if (d & 16) T1 = Tprev; else T0 = Tprev; }
            gb = max(gb, wbest);
            Lk = nL; Uk = nU;
            // the most any diagonal of the band can still score: slide to the end of q or s with no further difference
            // (2 min(qn, sn - k) + k peaks at k = sn - qn)
            const int32_t kp = min(max(sn - qn, Lk), Uk);
            const int32_t ub = 2 * min(qn, sn - kp) + kp - d6;
            if (Lk > Uk || ub <= gb || d >= qn + sn) st = GH_FIN;
        }
    }
    // totals
    for (int o = 16; o; o >>= 1) {
        cells += __shfl_down_sync(FULL, cells, o);
        produced += __shfl_down_sync(FULL, produced, o);
    }
    if (lane == 0) {
        if (cells) atomicAdd(ctr + 1, cells);
        if (produced) atomicAdd(ctr + 3, produced);
    }
    if (j == 0 && n_deferred && !in_list) atomicAdd(ctr + 2, n_deferred);
}
