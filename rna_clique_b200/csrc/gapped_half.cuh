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

#define GH_CELLS 3                     // adjacent diagonals per lane
#define GH_WIN (16 * GH_CELLS)         // diagonals per half-warp window
#define GH_KB0 (-(GH_WIN / 2))         // window start of a new extension: diagonal 0 in the middle
#define GH_THREADS 128

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

__global__ void __launch_bounds__(GH_THREADS, 4) k_gapped_half(
    const Ungapped* __restrict__ U, const uint64_t* __restrict__ keys, int64_t n, KeyFmt F,
    const uint32_t* __restrict__ heads, int64_t n_groups, PairView P, Gapped* __restrict__ out,
    uint32_t* __restrict__ deferred,
    unsigned long long* __restrict__ ctr /* [0]=next group [1]=cells [2]=retries [3]=n gapped [4]=#deferred groups */) {
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
    int32_t R0 = NEG, R1 = NEG, R2 = NEG;
    uint32_t M0 = 0, M1 = 0, M2 = 0, G0 = 0, G1 = 0, G2 = 0;
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
                    if (j == 0) deferred[atomicAdd(ctr + 4, 1ull)] = (uint32_t)h;
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
                if (p && j == 0) g = atomicAdd(ctr, 1ull);
                g = __shfl_sync(FULL, g, 0, 16);
                if (p && (int64_t)g >= n_groups) { st = GH_DONE; p = false; }
                if (p) {
                    h = heads[g];
                    e_ = h + 1;
                    const uint64_t key = keys[h];
                    while (e_ < n && keys[e_] == key) e_++;
                    gsize = (int32_t)(e_ - h);
                    ta = (int32_t)(key >> (F.tb_bits + 1));
                    tb = (int32_t)((key >> 1) & ((1u << F.tb_bits) - 1u));
                    minus = (int32_t)(key & 1);
                    la = __ldg(P.lenA + ta); lb = __ldg(P.lenB + tb);
                    if (la > RCQ_STAGE_MAX || lb > RCQ_STAGE_MAX) {
                        if (j == 0) deferred[atomicAdd(ctr + 4, 1ull)] = (uint32_t)h;
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
                    R0 = R1 = R2 = NEG; M0 = M1 = M2 = 0; G0 = G1 = G2 = 0;
                    const bool mine = (0 - GH_KB0) / GH_CELLS == j;     // diagonal 0 = cell (0 - kb) % 3 of lane (0 - kb) / 3
                    if (mine) {
                        if ((0 - GH_KB0) % GH_CELLS == 0) R0 = first;
                        else if ((0 - GH_KB0) % GH_CELLS == 1) R1 = first;
                        else R2 = first;
                    }
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
        bool act = st == GH_RUN;
        d += act ? 1 : 0;
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
                const int32_t r0 = __shfl_sync(FULL, R0, sl, 16), r1 = __shfl_sync(FULL, R1, sl, 16), r2 = __shfl_sync(FULL, R2, sl, 16);
                const uint32_t m0 = __shfl_sync(FULL, M0, sl, 16), m1 = __shfl_sync(FULL, M1, sl, 16), m2 = __shfl_sync(FULL, M2, sl, 16);
                const uint32_t g0 = __shfl_sync(FULL, G0, sl, 16), g1 = __shfl_sync(FULL, G1, sl, 16), g2 = __shfl_sync(FULL, G2, sl, 16);
                R0 = in ? r0 : NEG; R1 = in ? r1 : NEG; R2 = in ? r2 : NEG;
                M0 = m0; M1 = m1; M2 = m2; G0 = g0; G1 = g1; G2 = g2;
            }
        }
        const int32_t d6 = 6 * d;
        const int32_t xcut = tref - RCQ_XDROP_GAPPED2 + d6;       // alive iff 2 * row + k >= xcut
        // the two neighbour cells this lane does not hold
        int32_t Rl = __shfl_up_sync(FULL, R2, 1, 16), Rr = __shfl_down_sync(FULL, R0, 1, 16);
        const uint32_t Ml = __shfl_up_sync(FULL, M2, 1, 16), Gl = __shfl_up_sync(FULL, G2, 1, 16);
        const uint32_t Mr = __shfl_down_sync(FULL, M0, 1, 16), Gr = __shfl_down_sync(FULL, G0, 1, 16);
        if (j == 0) Rl = NEG;
        if (j == 15) Rr = NEG;
        const int32_t k0 = kb + GH_CELLS * j;
        SeqShared Qs, Ss;
        Qs.b = X.Q; Ss.b = X.S;

        int32_t nR[GH_CELLS], bi_[GH_CELLS], lim_[GH_CELLS];
        uint32_t nM[GH_CELLS], nG[GH_CELLS];
        bool alive[GH_CELLS], cont[GH_CELLS];
#pragma unroll
        for (int c = 0; c < GH_CELLS; c++) {
            const int32_t k = k0 + c;
            const int32_t left = c == 0 ? Rl : (c == 1 ? R0 : R1);
            const int32_t mid = c == 0 ? R0 : (c == 1 ? R1 : R2);
            const int32_t right = c == 0 ? R1 : (c == 1 ? R2 : Rr);
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
            const bool ok = act && bi >= 0 && 2 * bi + k >= xcut;
            const uint32_t sm = t == 2 ? (c == 0 ? M0 : (c == 1 ? M1 : M2)) : (t == 1 ? (c == 0 ? M1 : (c == 1 ? M2 : Mr)) : (c == 0 ? Ml : (c == 1 ? M0 : M1)));
            const uint32_t sg = t == 2 ? (c == 0 ? G0 : (c == 1 ? G1 : G2)) : (t == 1 ? (c == 0 ? G1 : (c == 1 ? G2 : Gr)) : (c == 0 ? Gl : (c == 1 ? G0 : G1)));
            const uint32_t op = 2u - (uint32_t)t;                 // 0 mismatch, 1 query gap, 2 subject gap
            const uint32_t gapw = ((sg & ~7u) | op) + (1u << 17) + (((sg & 7u) != op) ? 8u : 0u);
            nM[c] = sm + (t == 2 ? 1u : 0u);
            nG[c] = t == 2 ? (sg & ~7u) : gapw;
            const int32_t lim = rowmax - bi;
            // dead lanes compare the first words of the strands (in bounds) and drop the result
            const int32_t r16 = run16(Qs, ok ? X.qo + bi : 0, Ss, ok ? X.so + bi + k : 0);
            const int32_t z = (ok && lim > 0) ? min(r16, lim) : 0;
            nstep += ok ? (uint32_t)(z + 1) : 0u;
            alive[c] = ok; bi_[c] = bi; lim_[c] = lim;
            nR[c] = bi + z;
            cont[c] = ok && z == 16 && lim > 16;
        }
        // slides that survived their first 16 bases are finished by the half-warp together
#pragma unroll
        for (int c = 0; c < GH_CELLS; c++) {
            unsigned more = __ballot_sync(FULL, cont[c]);
            while (more) {
                const unsigned mh = (more >> hs) & 0xFFFFu;
                const int sl = mh ? __ffs(mh) - 1 : 0;
                const int32_t e2 = __shfl_sync(FULL, nR[c], sl, 16), l2 = __shfl_sync(FULL, lim_[c], sl, 16);
                const int32_t k2 = kb + GH_CELLS * sl + c;
                const int32_t extra = gh_coop_slide(X.Q, X.qo + e2, X.S, X.so + e2 + k2, mh ? l2 - 16 : 0, j, half);
                if (mh && j == sl) { nR[c] += extra; nstep += (uint32_t)extra; cont[c] = false; }
                more = __ballot_sync(FULL, cont[c]);
            }
        }
        int32_t wbest = NEG, ubmax = NEG, nL = 1 << 28, nU = -(1 << 28);
#pragma unroll
        for (int c = 0; c < GH_CELLS; c++) {
            const int32_t k = k0 + c;
            const bool ok = alive[c];
            const int32_t e = nR[c];
            if (ok && e > bi_[c]) nG[c] |= 4u;
            const int32_t sc = ok ? 2 * e + k - d6 : NEG;
            wbest = max(wbest, sc);
            nL = min(nL, ok ? k : (1 << 28)); nU = max(nU, ok ? k : -(1 << 28));
            ubmax = max(ubmax, ok ? 2 * min(qn, sn - k) + k - d6 : NEG);
            if (sc > bS) { bS = sc; bD = d; bK = k; bM = (int32_t)nM[c]; bG = (int32_t)nG[c]; }
            if (!ok) nR[c] = NEG;
        }
        if (act) {
            R0 = nR[0]; R1 = nR[1]; R2 = nR[2];
            M0 = nM[0]; M1 = nM[1]; M2 = nM[2];
            G0 = nG[0]; G1 = nG[1]; G2 = nG[2];
        }
#pragma unroll
        for (int o = 8; o; o >>= 1) {
            wbest = max(wbest, __shfl_xor_sync(FULL, wbest, o, 16));
            ubmax = max(ubmax, __shfl_xor_sync(FULL, ubmax, o, 16));
            nL = min(nL, __shfl_xor_sync(FULL, nL, o, 16));
            nU = max(nU, __shfl_xor_sync(FULL, nU, o, 16));
        }
        if (act) {
            Tprev = max(Tprev, wbest);
            if (j == (d & 15)) { if (d & 16) T1 = Tprev; else T0 = Tprev; }
            gb = max(gb, wbest);
            Lk = nL; Uk = nU;
            if (Lk > Uk || ubmax <= gb || d >= qn + sn) st = GH_FIN;
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
    if (j == 0 && n_deferred) atomicAdd(ctr + 2, n_deferred);
}
