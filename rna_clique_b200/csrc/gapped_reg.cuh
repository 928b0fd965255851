// gapped_reg.cuh — kernel (3b), main path: the greedy X-drop gapped extension of gapped_kernels.cuh (same
// recurrence, same results) with the wave rows in REGISTERS and several independent
// extension engines per warp.
//
// Why: the extension is bound by integer instruction issue.  A wave of the recurrence is narrow (1 diagonal at
// the start, about 2 x 19 + 1 once BLAST's X-drop window has opened), and a large part of the instructions of
// a wave does not depend on the number of cells a lane computes: the window test, the three reductions, the X-drop
// reference, the early-stop test, the state machine between extensions.  So
//   * an ENGINE is LANES lanes (16 or 8) that own one (transcript, transcript, strand) group; a warp holds
//     32 / LANES engines that run different groups through ONE instruction stream, which divides the fixed part
//     of a wave by the number of engines;
//   * every lane owns CELLS ADJACENT diagonals (LANES x CELLS = 48 or 64 per engine) in registers: furthest row R
//     and gap word G of the previous wave (the mismatch count is d - gaps).  The two neighbours a lane lacks come
//     from two shuffles, so a wave needs no shared-memory rows, no stores, no sentinels, no index arithmetic;
//   * the three candidate entries carry their tie rank in the two low bits (row << 2 | rank), so the entry and
//     the operation that made it come out of one three-way maximum (VIMNMX3); a candidate beyond a sequence end
//     exceeds the row limit by exactly one, so one compare sends the (rare) end-of-sequence case to the full
//     per-candidate clamps;
//   * the window follows the band cell by cell when indels make it drift; a band wider than the window, or
//     transcripts longer than the staged limit, defer the WHOLE group to the next pass (a wider window, finally
//     k_gapped with global rows of any width);
//   * only the TWO strands the running extension reads are staged in shared memory (the left part reads both
//     transcripts on the strands opposite to the hit, the right part on the hit's), copied by the whole warp;
//   * BLAST's X-drop reference (best score 19 waves back) lives in a 32-entry ring per engine in shared memory.
// The main loop is a state machine: the divergent part (fetch a group, pick the next ungapped HSP in canonical
// order, containment / duplicate checks, reduce the best end point, stage) runs for all engines of the warp
// together at the top of the loop with predicates; the wave step below it is common.
// EVERY warp primitive uses the FULL mask at a point all 32 lanes reach together: per-engine masks compile to
// MATCH.ANY + WARPSYNC.COLLECTIVE and leave the engines split into sub-warps for good (measured: 4x slower).
//
// Roofline: instruction issue (INT32); HBM traffic is the transcripts once per extension.  Work unit: extension
// steps (base comparisons), counted exactly as the oracle counts them.
#pragma once
#include "gapped_kernels.cuh"

enum { GR_FETCH = 0, GR_NEXT = 1, GR_RUN = 2, GR_FIN = 3, GR_DONE = 4 };

// every engine slides along one diagonal: number of matching base pairs from (xq, xs), at most lim (lim <= 0:
// nothing to do for this engine).  LANES lanes x 16 bases per round.
template <int LANES>
__device__ __forceinline__ int32_t gr_coop_slide(const uint32_t* Q, int32_t xq, const uint32_t* S, int32_t xs, int32_t lim,
                                                 int j, int hs) {
    constexpr unsigned EM = (1u << LANES) - 1u;
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
        const unsigned stop = (__ballot_sync(FULL, z < 16) >> hs) & EM;
        const int src = stop ? __ffs(stop) - 1 : 0;
        const int32_t zs = __shfl_sync(FULL, z, src, LANES);
        if (!done && stop) { result = base + 16 * src + zs; done = true; }
        base += 16 * LANES;
        if (!done && base >= lim) { result = lim; done = true; }
        if (__all_sync(FULL, done)) break;
    }
    return result;
}

struct GrExt {            // one extension of an engine (uniform over its lanes)
    const uint32_t *Q, *S;     // staged words of the two strands, already advanced to the word of the start base
    int32_t qo, so;            // offset of the start base inside that word (0 .. 15)
    int32_t qn, sn;
};

#define GR_ENGINE_WORDS (2 * RCQ_STAGE_W + 33)   // two staged strands + the X-drop ring (+1: bank spread)

// dynamic shared memory of one block: per engine the two strands of the running extension and the X-drop ring
template <int LANES, int THREADS>
constexpr size_t gr_smem_bytes() { return (size_t)(THREADS / LANES) * GR_ENGINE_WORDS * sizeof(uint32_t); }

// Groups come from `heads` (in_list == NULL) or from a list of group heads written by an earlier pass; groups given
// up (band wider than LANES x CELLS diagonals, transcript longer than the staged limit) are appended to out_list.
template <int LANES, int CELLS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_gapped_reg(
    const Ungapped* __restrict__ U, const uint64_t* __restrict__ keys, int64_t n, KeyFmt F,
    const uint32_t* __restrict__ heads, int64_t n_groups, PairView P, Gapped* __restrict__ out,
    const uint32_t* __restrict__ in_list, const unsigned long long* __restrict__ in_count, unsigned long long* __restrict__ in_next,
    uint32_t* __restrict__ out_list, unsigned long long* __restrict__ out_count,
    unsigned long long* __restrict__ ctr /* [1]=cells [2]=groups given up by the first pass [3]=n gapped */) {
    static_assert(LANES == 16 || LANES == 8, "engine width");
    constexpr unsigned EM = (1u << LANES) - 1u;
    constexpr int WIN = LANES * CELLS;           // diagonals per engine window
    constexpr int KB0 = -(WIN / 2);              // window start of a new extension: diagonal 0 in the middle
    constexpr int ENG_WARP = 32 / LANES;
    extern __shared__ uint32_t gr_smem[];
    const int lane = threadIdx.x & 31, j = lane & (LANES - 1);
    const int hs = lane & ~(LANES - 1);             // shift of this engine's bits in a ballot
    const int eng = threadIdx.x / LANES;            // engine of the block
    uint32_t* stage = gr_smem + (size_t)eng * GR_ENGINE_WORDS;
    int32_t* tring = (int32_t*)(stage + 2 * RCQ_STAGE_W);
    uint32_t* warp_stage0 = gr_smem + (size_t)((threadIdx.x >> 5) * ENG_WARP) * GR_ENGINE_WORDS;

    int st = GR_FETCH;
    // group
    int64_t h = 0, e_ = 0;
    int32_t gsize = 0, ta = 0, tb = 0, minus = 0, la = 0, lb = 0, count = 0, it = 0;
    int64_t oa = 0, ob = 0;               // first base of the two transcripts in their arenas
    Ungapped last; last.score = 0; last.diag = 0; last.a0 = 0; last.a1 = 0; last.i0 = 0; last.ta = last.tb = last.minus = 0;
    bool have_last = false, group_open = false;
    // current HSP / extension
    int32_t i0 = 0, j0 = 0, phase = 0;
    GrExt X; X.Q = stage; X.S = stage; X.qo = X.so = 0; X.qn = X.sn = 0;
    int32_t Lscore2 = 0, Lqi = 0, Lsj = 0, Ld = 0, Lgw = 0;
    // wave state
    int32_t d = 0, Lk = 0, Uk = 0, gb = 0, Tprev = 0, kb = KB0;
    int32_t R[CELLS];
    uint32_t G[CELLS];
#pragma unroll
    for (int c = 0; c < CELLS; c++) { R[c] = NEG; G[c] = 0; }
    int32_t bS = NEG, bD = 0, bK = 0, bG = 0;
    bool overflow = false;
    uint32_t nstep = 0;                     // extension steps of the open group (this lane's share)
    unsigned long long cells = 0, produced = 0, n_deferred = 0;

    while (true) {
        // ------------------------------------------------------------------ set-up passes (all engines together)
        while (__any_sync(FULL, st != GR_RUN && st != GR_DONE)) {
            bool begin = false;                       // this engine starts an extension at the end of the pass
            int32_t bqrel = 0, bqn = 0, bsrel = 0, bsn = 0;
            // ---- (1) an extension has ended
            if (__any_sync(FULL, st == GR_FIN)) {
                bool p = st == GR_FIN;
                if (p && overflow) {
                    // the band outgrew the window: the whole group goes to the next pass
                    if (j == 0) out_list[atomicAdd(out_count, 1ull)] = (uint32_t)h;
                    n_deferred++;
                    group_open = false; nstep = 0; overflow = false;
                    st = GR_FETCH; p = false;
                }
                // best end point of the extension: score desc, then d asc, then k asc
                uint64_t key = 0;
                if (p && bS > NEG)
                    key = ((uint64_t)(uint32_t)(bS + (1 << 20)) << 42) | ((uint64_t)((1 << 21) - 1 - bD) << 21) |
                          (uint64_t)((1 << 21) - 1 - (bK + (1 << 20)));
                uint64_t best = key;
#pragma unroll
                for (int o = LANES / 2; o; o >>= 1) { const uint64_t x = __shfl_xor_sync(FULL, best, o, LANES); if (x > best) best = x; }
                const unsigned who = (__ballot_sync(FULL, key == best) >> hs) & EM;
                const int src = who ? __ffs(who) - 1 : 0;
                const int32_t rS = __shfl_sync(FULL, bS, src, LANES), rD = __shfl_sync(FULL, bD, src, LANES),
                              rK = __shfl_sync(FULL, bK, src, LANES), rG = __shfl_sync(FULL, bG, src, LANES);
                const int32_t rI = (rS - rK + 6 * rD) >> 1;          // score = 2 row + k - 6 d
                bool dup = false;
                Gapped gp;
                gp.ta = ta; gp.tb = tb; gp.minus = minus;
                gp.a0 = i0 + 1 - Lqi; gp.a1 = i0 + 1 + rI;
                gp.b0 = j0 + 1 - Lsj; gp.b1 = j0 + 1 + rI + rK;
                gp.score2 = Lscore2 + rS;
                gp.gaps = (Lgw >> 17) + (rG >> 17);
                gp.mism = Ld + rD - gp.gaps;                          // every wave adds one difference: a mismatch or one gap base
                gp.gapopen = ((Lgw >> 3) & 0x3fff) + ((rG >> 3) & 0x3fff);
                gp.nident = ((gp.a1 - gp.a0) + (gp.b1 - gp.b0) - gp.gaps) / 2 - gp.mism;
                const bool fin_r = p && phase == 1;
                if (fin_r) {
                    for (int32_t c0 = j; c0 < count; c0 += LANES) {
                        const Gapped o2 = out[h + c0];
                        dup = dup || (o2.a0 == gp.a0 && o2.a1 == gp.a1 && o2.b0 == gp.b0 && o2.b1 == gp.b1);
                    }
                }
                const bool any_dup = (__ballot_sync(FULL, dup) >> hs) & EM;
                if (fin_r) {
                    if (!any_dup) {
                        if (j == 0) out[h + count] = gp;
                        count++;
                    }
                    st = GR_NEXT;
                } else if (p) {           // the left part is done: keep it, start the right part
                    Lscore2 = rS; Lqi = rI; Lsj = rI + rK; Ld = rD; Lgw = rG;
                    phase = 1;
                    begin = true;
                    bqrel = i0 + 1; bqn = la - i0 - 1; bsrel = j0 + 1; bsn = lb - j0 - 1;
                }
                __syncwarp();             // out[] written by lane 0 of an engine is read by its other lanes below
            }
            // ---- (2) take the next group
            if (__any_sync(FULL, st == GR_FETCH)) {
                bool p = st == GR_FETCH;
                if (p && group_open) {        // the finished group's work counts; a deferred one was dropped above
                    cells += nstep;
                    produced += (j == 0) ? (unsigned long long)count : 0ull;
                    nstep = 0;
                    group_open = false;
                }
                unsigned long long g = 0;
                if (p && j == 0) g = atomicAdd(in_next, 1ull);
                g = __shfl_sync(FULL, g, 0, LANES);
                if (p && (int64_t)g >= (in_list ? (int64_t)*in_count : n_groups)) { st = GR_DONE; p = false; }
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
                        p = false;            // stays in GR_FETCH: next pass takes another group
                    } else {
                        oa = __ldg(P.offA + ta); ob = __ldg(P.offB + tb);
                    }
                }
                if (p) {
                    count = 0; it = 0; have_last = false; group_open = true;
                    st = GR_NEXT;
                }
            }
            // ---- (3) the next ungapped HSP of the group, strictly after `last` in canonical order
            if (__any_sync(FULL, st == GR_NEXT)) {
                bool p = st == GR_NEXT;
                if (p && (it >= gsize || count >= RCQ_MAX_GAPPED_PER_GROUP)) { st = GR_FETCH; p = false; }
                if (p) it++;
                Ungapped bestu; bool found = false;
                bestu.score = 0; bestu.diag = 0; bestu.a0 = 0; bestu.a1 = 0; bestu.i0 = 0;
                if (p) {
                    for (int64_t x = h + j; x < e_; x += LANES) {
                        const Ungapped u = U[x];
                        if (u.score < 0) continue;
                        if (have_last && !ung_before(last, u)) continue;
                        if (!found || ung_before(u, bestu)) { bestu = u; found = true; }
                    }
                }
#pragma unroll
                for (int o = LANES / 2; o; o >>= 1) {
                    Ungapped v;
                    v.score = __shfl_xor_sync(FULL, bestu.score, o, LANES); v.diag = __shfl_xor_sync(FULL, bestu.diag, o, LANES);
                    v.a0 = __shfl_xor_sync(FULL, bestu.a0, o, LANES); v.a1 = __shfl_xor_sync(FULL, bestu.a1, o, LANES);
                    v.i0 = __shfl_xor_sync(FULL, bestu.i0, o, LANES);
                    const bool f2 = __shfl_xor_sync(FULL, (int)found, o, LANES);
                    if (f2 && (!found || ung_before(v, bestu))) {
                        bestu.score = v.score; bestu.diag = v.diag; bestu.a0 = v.a0; bestu.a1 = v.a1; bestu.i0 = v.i0; found = true;
                    }
                }
                if (p && !found) { st = GR_FETCH; p = false; }
                bool contained = false;
                if (p) {
                    last = bestu; have_last = true;
                    const int32_t ub0 = bestu.a0 + bestu.diag, ub1 = bestu.a1 + bestu.diag;
                    for (int32_t c0 = j; c0 < count; c0 += LANES) {
                        const Gapped gq = out[h + c0];
                        contained = contained || (gq.a0 <= bestu.a0 && bestu.a1 <= gq.a1 && gq.b0 <= ub0 && ub1 <= gq.b1);
                    }
                }
                const bool any_contained = (__ballot_sync(FULL, contained) >> hs) & EM;
                if (p && !any_contained) {
                    i0 = bestu.i0; j0 = i0 + bestu.diag;
                    phase = 0;
                    // left part: the anchor pair first, walking backwards = forwards on the opposite strands
                    begin = true;
                    bqrel = la - 1 - i0; bqn = i0 + 1; bsrel = lb - 1 - j0; bsn = j0 + 1;
                }
            }
            // ---- (4) start the extensions asked for above: stage, wave 0 and the register rows
            if (__any_sync(FULL, begin)) {
                // the two strands this extension reads (left part: A and B on the strands opposite to the hit, so that
                // walking backwards is walking forwards), copied by the whole warp for each engine in turn
                unsigned need = __ballot_sync(FULL, begin && j == 0);
                while (need) {
                    const int ld = __ffs(need) - 1;
                    need &= need - 1;
                    const int64_t oa_ = __shfl_sync(FULL, oa, ld), ob_ = __shfl_sync(FULL, ob, ld);
                    const int32_t la_ = __shfl_sync(FULL, la, ld), lb_ = __shfl_sync(FULL, lb, ld);
                    const int32_t mi_ = __shfl_sync(FULL, minus, ld), ph_ = __shfl_sync(FULL, phase, ld);
                    uint32_t* dst = warp_stage0 + (size_t)(ld / LANES) * GR_ENGINE_WORDS;
                    const uint32_t* ga = (ph_ ? P.Af.bits : P.Ar.bits) + (oa_ >> 4);
                    const uint32_t* gs = ((ph_ != 0) == (mi_ != 0) ? P.Br.bits : P.Bf.bits) + (ob_ >> 4);
                    const int32_t wa = (la_ + 15) / 16 + 2, wb = (lb_ + 15) / 16 + 2;
                    for (int32_t w = lane; w < wa; w += 32) dst[w] = __ldg(ga + w);
                    for (int32_t w = lane; w < wb; w += 32) dst[RCQ_STAGE_W + w] = __ldg(gs + w);
                }
                __syncwarp();             // the staged words are read by other lanes
                if (begin) {
                    X.Q = stage + (bqrel >> 4); X.qo = bqrel & 15; X.qn = bqn;
                    X.S = stage + RCQ_STAGE_W + (bsrel >> 4); X.so = bsrel & 15; X.sn = bsn;
                }
                const int32_t first = gr_coop_slide<LANES>(X.Q, X.qo, X.S, X.so, begin ? min(X.qn, X.sn) : 0, j, hs);
                if (begin) {
                    if (j == 0) nstep += (uint32_t)(first + 1);
                    kb = KB0;
                    const bool mine = (0 - KB0) / CELLS == j;     // diagonal 0 = cell (0 - kb) % CELLS of lane (0 - kb) / CELLS
#pragma unroll
                    for (int c = 0; c < CELLS; c++) { R[c] = (mine && c == (0 - KB0) % CELLS) ? first : NEG; G[c] = 0; }
                    bS = mine ? 2 * first : NEG; bD = 0; bK = 0; bG = 0;
                    gb = 2 * first; Tprev = gb;
                    if (j == 0) tring[0] = gb;
                    Lk = 0; Uk = 0; d = 0;
                    overflow = false;
                    st = (X.qn + X.sn >= 1) ? GR_RUN : GR_FIN;
                }
                __syncwarp();
            }
        }
        if (__all_sync(FULL, st == GR_DONE)) break;

        // ---------------------------------------------------------------------- one wave, all engines
        const bool act0 = st == GR_RUN;
        d += act0 ? 1 : 0;
        const int32_t qn = X.qn, sn = X.sn;
        int32_t tref = tring[(d - RCQ_XDROP_OFFSET) & 31];
        if (d < RCQ_XDROP_OFFSET) tref = 0;
        const int32_t klo = max(Lk - 1, -qn), khi = min(Uk + 1, sn);
        bool act = act0;
        // keep the band inside the window: move the window cell by cell, or give the group up when the band is wider
        {
            const bool out_of_window = act && (klo < kb || khi > kb + WIN - 1);
            if (__any_sync(FULL, out_of_window)) {
                int32_t diff = 0;                // cells to move: new cell i takes the content of old cell i + diff
                if (out_of_window) {
                    const int32_t width = khi - klo + 1;
                    if (width > WIN) {
                        overflow = true; st = GR_FIN; act = false;
                    } else {
                        diff = klo - (WIN - width) / 2 - kb;
                        kb += diff;
                    }
                }
                const int32_t sl = diff / CELLS, sc = diff % CELLS;      // whole lanes, then single cells of the same sign
                for (int32_t t = 0; __any_sync(FULL, t < sc); t++) {     // up: cell c takes cell c + 1
                    const int32_t cr = __shfl_down_sync(FULL, R[0], 1, LANES);
                    const uint32_t cg = __shfl_down_sync(FULL, G[0], 1, LANES);
                    if (t < sc) {
#pragma unroll
                        for (int c = 0; c + 1 < CELLS; c++) { R[c] = R[c + 1]; G[c] = G[c + 1]; }
                        R[CELLS - 1] = (j == LANES - 1) ? NEG : cr;
                        G[CELLS - 1] = cg;
                    }
                }
                for (int32_t t = 0; __any_sync(FULL, t < -sc); t++) {    // down: cell c takes cell c - 1
                    const int32_t cr = __shfl_up_sync(FULL, R[CELLS - 1], 1, LANES);
                    const uint32_t cg = __shfl_up_sync(FULL, G[CELLS - 1], 1, LANES);
                    if (t < -sc) {
#pragma unroll
                        for (int c = CELLS - 1; c > 0; c--) { R[c] = R[c - 1]; G[c] = G[c - 1]; }
                        R[0] = (j == 0) ? NEG : cr;
                        G[0] = cg;
                    }
                }
                if (__any_sync(FULL, sl != 0)) {
                    const int srcl = j + sl;
                    const bool in = srcl >= 0 && srcl < LANES;
                    const int sidx = in ? srcl : j;
#pragma unroll
                    for (int c = 0; c < CELLS; c++) {
                        const int32_t r = __shfl_sync(FULL, R[c], sidx, LANES);
                        G[c] = __shfl_sync(FULL, G[c], sidx, LANES);
                        R[c] = in ? r : NEG;
                    }
                }
            }
        }
        const int32_t d6 = 6 * d;
        // alive iff 2 * row + k >= xcut (dead predecessors are far below any cut); an engine that is not running gets
        // a cut nothing passes
        const int32_t xcut = act ? tref - RCQ_XDROP_GAPPED2 + d6 : 0x7fffffff;
        // predecessor rows / gap words of cells -1 .. CELLS: the two outer ones come from the neighbour lanes
        int32_t Rx[CELLS + 2];
        uint32_t Gx[CELLS + 2];
        Rx[0] = __shfl_up_sync(FULL, R[CELLS - 1], 1, LANES); Rx[CELLS + 1] = __shfl_down_sync(FULL, R[0], 1, LANES);
        Gx[0] = __shfl_up_sync(FULL, G[CELLS - 1], 1, LANES); Gx[CELLS + 1] = __shfl_down_sync(FULL, G[0], 1, LANES);
        if (j == 0) Rx[0] = NEG;
        if (j == LANES - 1) Rx[CELLS + 1] = NEG;
#pragma unroll
        for (int c = 0; c < CELLS; c++) { Rx[c + 1] = R[c]; Gx[c + 1] = G[c]; }
        const int32_t k0 = kb + CELLS * j;
        SeqShared Qs, Ss;
        Qs.b = X.Q; Ss.b = X.S;

        int32_t zsum = 0;
        uint32_t alive = 0;                       // bit c: cell c of this lane is inside the X-drop band
        // the first cell of the lane whose slide ran past its 16 bases: finished below by the whole engine
        bool has_cont = false;
        int32_t ce = 0, cl = 0, ck = 0, cidx = 0;
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
            const uint32_t sg = mis ? Gx[c + 1] : (qgap ? Gx[c + 2] : Gx[c]);
            const uint32_t op = 2u - (uint32_t)t;                 // 0 mismatch, 1 query gap, 2 subject gap
            const uint32_t base = sg & ~7u;
            // a gap base: one more gap, and one more opening unless the same kind of gap ended right here
            const uint32_t gapw = (base | op) + (1u << 17) + ((((sg ^ op) & 7u) + 7u) & 8u);
            const int32_t lim = rowmax - bi;
            // dead cells compare in-bounds words of the strands and drop the result
            const int32_t bs = max(bi, 0);
            const int32_t r16 = run16(Qs, X.qo + bs, Ss, max(X.so + bs + k, 0));
            int32_t z = ok ? min(r16, max(lim, 0)) : 0;
            if (z == 16 && lim > 16) {
                if (!has_cont) {
                    has_cont = true; ce = bi + 16; cl = lim - 16; ck = k; cidx = c;
                } else {
                    // a second long slide in the same lane (rare): finished by the lane alone
                    int32_t rem = lim - 16;
                    while (rem > 0) {
                        const int32_t r = min(run16(Qs, X.qo + bi + z, Ss, X.so + bi + z + k), rem);
                        z += r; rem -= r;
                        if (r < 16) break;
                    }
                }
            }
            zsum += z;
            R[c] = ok ? bi + z : NEG;
            G[c] = (mis ? base : gapw) | (z > 0 ? 4u : 0u);
            alive |= ok ? (1u << c) : 0u;
        }
        // slides that survived their first 16 bases are finished by the engine together
        if (__any_sync(FULL, has_cont)) {
            bool pend = has_cont;
            while (true) {
                const unsigned wb = __ballot_sync(FULL, pend);
                if (!wb) break;
                const unsigned w = (wb >> hs) & EM;
                const int sidx = w ? __ffs(w) - 1 : 0;
                const int32_t e2 = __shfl_sync(FULL, ce, sidx, LANES), l2 = __shfl_sync(FULL, cl, sidx, LANES),
                              k2 = __shfl_sync(FULL, ck, sidx, LANES);
                const int32_t extra = gr_coop_slide<LANES>(X.Q, X.qo + e2, X.S, X.so + e2 + k2, w ? l2 : 0, j, hs);
                if (w && j == sidx) {
#pragma unroll
                    for (int c = 0; c < CELLS; c++) R[c] += (c == cidx) ? extra : 0;
                    zsum += extra;
                    pend = false;
                }
            }
        }
        // best cell of the lane in this wave: score first, then the lowest diagonal (dead cells: NEG far below)
        int32_t lkey = NEG;
#pragma unroll
        for (int c = 0; c < CELLS; c++) {
            const int32_t sc = R[c] >= 0 ? 2 * R[c] + (k0 + c) - d6 : (NEG >> 4);
            lkey = max(lkey, sc * 16 + (15 - c));
        }
        int32_t wbest = lkey >> 4;
        if (alive && wbest > bS) {                // strictly better than every earlier wave: (d, k) ties keep the first
            const int cb = 15 - (lkey & 15);
            bS = wbest; bD = d; bK = k0 + cb;
            uint32_t g = G[0];
#pragma unroll
            for (int c = 1; c < CELLS; c++) g = (c == cb) ? G[c] : g;
            bG = (int32_t)g;
        }
        nstep += (uint32_t)(zsum + __popc(alive));       // one step per base slid over plus one per live cell
        // live range of the band, both ends in one packed 16-bit maximum (|k| <= RCQ_STAGE_MAX): high = Uk, low = -Lk
        uint32_t rng = alive ? (((uint32_t)(uint16_t)(int16_t)(k0 + 31 - __clz(alive)) << 16) | (uint32_t)(uint16_t)(int16_t)(-(k0 + __ffs(alive) - 1)))
                             : 0x8AD08AD0u;       // (-30000, -30000): no live cell
#pragma unroll
        for (int o = LANES / 2; o; o >>= 1) {
            wbest = max(wbest, __shfl_xor_sync(FULL, wbest, o, LANES));
            rng = __vmaxs2(rng, __shfl_xor_sync(FULL, rng, o, LANES));
        }
        const int32_t nU = (int32_t)(int16_t)(rng >> 16), nL = -(int32_t)(int16_t)(rng & 0xFFFFu);
        if (act) {
            Tprev = max(Tprev, wbest);
            if (j == 0) tring[d & 31] = Tprev;
            gb = max(gb, wbest);
            Lk = nL; Uk = nU;
            // the most any diagonal of the band can still score: slide to the end of q or s with no further difference
            // (2 min(qn, sn - k) + k peaks at k = sn - qn)
            const int32_t kp = min(max(sn - qn, Lk), Uk);
            const int32_t ub = 2 * min(qn, sn - kp) + kp - d6;
            if (Lk > Uk || ub <= gb || d >= qn + sn) st = GR_FIN;
        }
        __syncwarp();                             // the ring entry is read 19 waves later by the other lanes
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
