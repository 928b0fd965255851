/*
 * rcq_oracle.c — CPU ORACLE (test infrastructure, NOT the product).
 *
 * Plain-C restatement of the per-pair search of RNA-clique's phase 1:
 *   reference find_homologs.py:82-133 (gene_matches -> blastn megablast, both directions),
 *   find_homologs.py:135-164 (highest_bitscores), find_homologs.py:215-302 (get_match_table).
 * The alignment arithmetic of the reference lives in NCBI BLAST+ `blastn` (third-party, version
 * unpinned: whatever `apt install ncbi-blast+` gives, reference tests/test_install/ubuntu.sh:18),
 * reached through simple-blast>=0.7.9 (pyproject.toml:33).  Neither is in /root/reference nor
 * installable here, and the reference has no golden vectors for this boundary, so for the
 * alignment stage this oracle is **PARITY UNPINNED**: it restates the published megablast
 * algorithm (word 28 exact seeds, +1/-2 scoring, linear gap cost 2.5, X-drop 20/100 bits,
 * Karlin-Altschul lambda=1.28 K=0.46 alpha=1.5 beta=-2, e-value cut, best hit per query gene) in
 * the canonical, order-independent form written down in DESIGN.md §3.  The table-assembly
 * stage (select_topn / match_table below) IS pinned: tests/golden/make_golden.py runs the
 * reference's own HomologFinder.get_match_table on synthetic hit frames and the fixtures are
 * checked against this file.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file.  Everything here is simple and slow on purpose (per-base loops, qsort).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define WORD 28
#define XU 10       /* ungapped X-drop, raw score units */
#define XG2 108     /* gapped X-drop, doubled score units */
#define M2 2
#define X2 (-4)
#define G2 (-5)
#define MAXG 32     /* gapped HSPs kept per (tx1, tx2, strand) group */
#define NEG (-(1 << 29))
#define KA_LAMBDA 1.28
#define KA_K 0.46
#define KA_ALPHA 1.5
#define KA_BETA (-2.0)

typedef struct {
    int32_t n_seq;
    int64_t total;
    int64_t* off;     /* n_seq+1 offsets into codes */
    uint8_t* fwd;     /* 0..3 = ACGT, 4 = anything else */
    uint8_t* rc;      /* per-transcript reverse complement, same offsets */
    int32_t *gene, *iso;
    uint8_t* dust;    /* 1 = base inside a symmetric-DUST perfect interval (seeding only), fwd coordinates */
    uint8_t* dust_rc; /* the same mask in reverse-complement coordinates */
} orc_sample;

/* ---------- symmetric DUST (Morgulis, Gertz, Schaffer, Agarwala 2006), blastn's default query filter
 * "-dust 20 64 1": level 20 (score threshold 2.0), window 64, linker 1.  Stated by its definition, not by the
 * streaming algorithm: over the triplets t_0 .. t_{m-1} of a maximal run of ACGT bases, an interval of
 * l >= 2 consecutive triplets has score r / (l - 1), r = sum over triplet values of c (c - 1) / 2; it is
 * PERFECT when it spans at most 62 triplets (64 bases), its score exceeds 2 and no interval inside it has a
 * higher score.  The mask is the union of the bases of all perfect intervals; linker 1 only joins masked
 * ranges that already touch, so it does not change the mask.  Soft masking: the mask keeps seeds out
 * (find_ungapped), extensions never look at it.  Pure function of the sequence, symmetric under reverse
 * complement.  Cost: a 62-step suffix scan per position; the full triangle only where a suffix scores > 2. */
#define DUST_MAXT 62
static int g_dust_on = 1;
void orc_set_dust(int on) { g_dust_on = on; }
static void dust_mask(const uint8_t* s, int32_t n, uint8_t* out) {
    memset(out, 0, n > 0 ? n : 0);
    if (!g_dust_on || n < 3) return;
    int32_t m = n - 2;
    int16_t* trip = malloc(sizeof(int16_t) * m);
    for (int32_t i = 0; i < m; i++)
        trip[i] = (s[i] > 3 || s[i + 1] > 3 || s[i + 2] > 3) ? -1 : (int16_t)(s[i] << 4 | s[i + 1] << 2 | s[i + 2]);
    uint8_t c[64];
    for (int32_t e = 1; e < m; e++) {
        if (trip[e] < 0) continue;
        /* suffixes [i, e]: does any score more than 2? */
        memset(c, 0, 64);
        int32_t r = 0, lo = -1;
        for (int32_t i = e; i >= 0 && e - i < DUST_MAXT && trip[i] >= 0; i--) {
            r += c[trip[i]]++;
            if (r > 2 * (e - i)) lo = i;       /* l - 1 = e - i */
        }
        if (lo < 0) continue;
        /* rows i = e-1 .. lo: best score of any interval inside [i, e] (bests_r / bests_l), and whether the
         * suffix [i, e] itself reaches it */
        int32_t best_r = 0, best_l = 1;
        for (int32_t i = e - 1; i >= lo; i--) {
            memset(c, 0, 64);
            r = 0;
            for (int32_t j = i; j <= e; j++) {
                r += c[trip[j]]++;
                if (j > i && r * best_l > best_r * (j - i)) { best_r = r; best_l = j - i; }
            }
            if (r > 2 * (e - i) && r * best_l >= best_r * (e - i))
                for (int32_t b = i; b <= e + 2; b++) out[b] = 1;
        }
    }
    free(trip);
}

static uint8_t code_of(uint8_t c) {
    c &= 0xDF;
    return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : (c == 'T' || c == 'U') ? 3 : 4;
}

orc_sample* orc_sample_new(const uint8_t* ascii, const int64_t* offsets, const int32_t* gene,
                           const int32_t* iso, int32_t n_seq) {
    orc_sample* s = calloc(1, sizeof *s);
    s->n_seq = n_seq;
    s->off = malloc(sizeof(int64_t) * (n_seq + 1));
    s->gene = malloc(sizeof(int32_t) * (n_seq + 1));
    s->iso = malloc(sizeof(int32_t) * (n_seq + 1));
    int64_t base = n_seq ? offsets[0] : 0;
    for (int t = 0; t <= n_seq; t++) s->off[t] = n_seq ? offsets[t] - base : 0;
    s->total = s->off[n_seq];
    s->fwd = malloc(s->total + 1);
    s->rc = malloc(s->total + 1);
    s->dust = malloc(s->total + 1);
    s->dust_rc = malloc(s->total + 1);
    for (int t = 0; t < n_seq; t++) {
        s->gene[t] = gene[t];
        s->iso[t] = iso[t];
        int64_t o = s->off[t], n = s->off[t + 1] - o;
        for (int64_t i = 0; i < n; i++) s->fwd[o + i] = code_of(ascii[base + o + i]);
        for (int64_t i = 0; i < n; i++) {
            uint8_t c = s->fwd[o + n - 1 - i];
            s->rc[o + i] = c > 3 ? 4 : (uint8_t)(3 - c);
        }
        dust_mask(s->fwd + o, (int32_t)n, s->dust + o);
        for (int64_t i = 0; i < n; i++) s->dust_rc[o + i] = s->dust[o + n - 1 - i];
    }
    return s;
}
/* test hook: the DUST mask of the whole sample, one byte per base (fwd coordinates) */
void orc_sample_dust(const orc_sample* s, uint8_t* out) { memcpy(out, s->dust, s->total); }
void orc_sample_free(orc_sample* s) {
    if (!s) return;
    free(s->off); free(s->fwd); free(s->rc); free(s->gene); free(s->iso); free(s->dust); free(s->dust_rc); free(s);
}

static inline int is_match(uint8_t a, uint8_t b) { return a == b && a < 4; }

/* ---------- Karlin-Altschul statistics (BLAST_ComputeLengthAdjustment, blastn 1/-2) ---------- */
int32_t orc_length_adjustment(int64_t query_length, int64_t db_length, int64_t db_num_seqs) {
    const double K = KA_K, logK = log(KA_K), adl = KA_ALPHA / KA_LAMBDA, beta = KA_BETA;
    double m = (double)query_length, n = (double)db_length, N = (double)db_num_seqs;
    double ell, ss, ell_bar, ell_min = 0, ell_max, ell_next = 0;
    int converged = 0;
    double a = N, mb = m * N + n, c = n * m - (m > n ? m : n) / K;
    if (c < 0) return 0;
    ell_max = 2 * c / (mb + sqrt(mb * mb - 4 * a * c));
    for (int i = 1; i <= 20; i++) {
        ell = ell_next;
        ss = (m - ell) * (n - N * ell);
        ell_bar = adl * (logK + log(ss)) + beta;
        if (ell_bar >= ell) {
            ell_min = ell;
            if (ell_bar - ell_min <= 1.0) { converged = 1; break; }
            if (ell_min == ell_max) break;
        } else {
            ell_max = ell;
        }
        if (ell_min <= ell_bar && ell_bar <= ell_max) ell_next = ell_bar;
        else ell_next = (i == 1) ? ell_max : (ell_min + ell_max) / 2;
    }
    int32_t la = (int32_t)ell_min;
    if (converged) {
        ell = ceil(ell_min);
        if (ell <= ell_max) {
            ss = (m - ell) * (n - N * ell);
            if (adl * (logK + log(ss)) + beta >= ell) la = (int32_t)ell;
        }
    }
    return la;
}
double orc_evalue(int32_t score, int64_t m, int64_t n, int64_t nseq) {
    int32_t la = orc_length_adjustment(m, n, nseq);
    double em = (double)(m - la), en = (double)(n - nseq * (int64_t)la);
    if (em < 1) em = 1;
    if (en < 1) en = 1;
    return KA_K * em * en * exp(-KA_LAMBDA * (double)score);
}
double orc_bitscore(int32_t score) { return (KA_LAMBDA * score - log(KA_K)) / log(2.0); }

/* ------------------------------- HSP records ------------------------------------------ */
typedef struct {
    int32_t ta, tb, minus;     /* transcript in sample 1, transcript in sample 2, strand of 2 */
    int32_t diag;              /* j' - i, j' on the given strand of tb */
    int32_t a0, a1;            /* half-open range on ta (plus strand) */
    int32_t score;             /* ungapped: raw score */
} ungapped_t;

typedef struct {
    int32_t ta, tb, minus;
    int32_t a0, a1, b0, b1;    /* half-open; b on the given strand of tb */
    int32_t score2;            /* doubled score */
    int32_t mism, gaps, gapopen, nident;
} gapped_t;

typedef struct { int32_t score2, qi, sj, mism, gaps, gapopen; } ext_t;

/* X-drop ungapped extension along one diagonal from an exact run [rs, re) (positions on a). */
static ungapped_t ungapped_extend(const uint8_t* a, int32_t la, const uint8_t* b, int32_t lb,
                                  int32_t rs, int32_t re, int32_t d) {
    int32_t sum = 0, best = 0, ext = 0;
    for (int32_t k = 0; re + k < la && re + d + k < lb; k++) {
        sum += is_match(a[re + k], b[re + d + k]) ? 1 : -2;
        if (sum > best) { best = sum; ext = k + 1; }
        else if (best - sum > XU) break;
    }
    ungapped_t u;
    u.a1 = re + ext;
    u.score = (re - rs) + best;
    sum = 0; best = 0; ext = 0;
    for (int32_t k = 0; rs - 1 - k >= 0 && rs - 1 - k + d >= 0; k++) {
        sum += is_match(a[rs - 1 - k], b[rs - 1 - k + d]) ? 1 : -2;
        if (sum > best) { best = sum; ext = k + 1; }
        else if (best - sum > XU) break;
    }
    u.a0 = rs - ext;
    u.score += best;
    u.diag = d;
    return u;
}

/* Start point of the gapped extension: middle of the longest exact run inside the HSP
 * (earliest run on ties). */
static int32_t gapped_start(const uint8_t* a, const uint8_t* b, const ungapped_t* u) {
    int32_t best_len = 0, best_s = u->a0, run = 0;
    for (int32_t i = u->a0; i < u->a1; i++) {
        if (is_match(a[i], b[i + u->diag])) {
            run++;
            if (run > best_len) { best_len = run; best_s = i - run + 1; }
        } else run = 0;
    }
    return best_s + best_len / 2;
}

/* X-drop DP with linear gap cost, anti-diagonal order.
 * q[0], q[qstep], ... (qn bases) against s[0], s[sstep], ... (sn bases).
 * Cell (i, j) = i bases of q and j bases of s consumed; t = i + j; k = j - i.
 * A cell is dead when its score < (best over anti-diagonals < t) - XG2.
 * Tie rule inside a cell: diagonal move, then U (consume q), then L (consume s).
 * Best end point: highest score, then smallest t, then smallest k. */
static int64_t g_dp_cells = 0;
static ext_t xdrop_extend(const uint8_t* q, int32_t qn, int qstep, const uint8_t* s, int32_t sn,
                          int sstep) {
    int32_t nd = qn + sn + 3;
    int32_t* H = malloc(sizeof(int32_t) * nd);
    int32_t* MM = calloc(nd, sizeof(int32_t));
    int32_t* GP = calloc(nd, sizeof(int32_t));
    int32_t* GO = calloc(nd, sizeof(int32_t));
    uint8_t* OP = calloc(nd, 1); /* 0 = D, 1 = U, 2 = L */
    for (int32_t i = 0; i < nd; i++) H[i] = NEG;
#define IX(k) ((k) + qn + 1)
    ext_t best = {0, 0, 0, 0, 0, 0};
    H[IX(0)] = 0;
    int32_t gbest = 0;
    /* live ranges (diagonals) of the two previous anti-diagonals; empty when lo > hi */
    int32_t lo1 = 0, hi1 = 0, lo2 = 1, hi2 = 0;
    for (int32_t t = 1; t <= qn + sn; t++) {
        int32_t lo, hi;
        if (lo1 > hi1 && lo2 > hi2) break;
        if (lo1 > hi1) { lo = lo2; hi = hi2; }
        else if (lo2 > hi2) { lo = lo1; hi = hi1; }
        else { lo = lo1 < lo2 ? lo1 : lo2; hi = hi1 > hi2 ? hi1 : hi2; }
        lo -= 1; hi += 1;
        if (lo < -qn) lo = -qn;
        if (hi > sn) hi = sn;
        int32_t nlo = 1, nhi = 0, tmax = NEG;
        for (int32_t k = lo; k <= hi; k++) {
            if (((t - k) & 1) != 0) continue;
            int32_t i = (t - k) / 2, j = (t + k) / 2;
            if (i < 0 || j < 0 || i > qn || j > sn) { continue; }
            g_dp_cells++;
            int32_t h = NEG, mm = 0, gp = 0, go = 0; uint8_t op = 0;
            if (i >= 1 && j >= 1 && H[IX(k)] > NEG) {
                int m = is_match(q[(int64_t)(i - 1) * qstep], s[(int64_t)(j - 1) * sstep]);
                h = H[IX(k)] + (m ? M2 : X2);
                mm = MM[IX(k)] + (m ? 0 : 1); gp = GP[IX(k)]; go = GO[IX(k)]; op = 0;
            }
            if (i >= 1 && H[IX(k + 1)] > NEG && H[IX(k + 1)] + G2 > h) {
                h = H[IX(k + 1)] + G2;
                mm = MM[IX(k + 1)]; gp = GP[IX(k + 1)] + 1;
                go = GO[IX(k + 1)] + (OP[IX(k + 1)] != 1); op = 1;
            }
            if (j >= 1 && H[IX(k - 1)] > NEG && H[IX(k - 1)] + G2 > h) {
                h = H[IX(k - 1)] + G2;
                mm = MM[IX(k - 1)]; gp = GP[IX(k - 1)] + 1;
                go = GO[IX(k - 1)] + (OP[IX(k - 1)] != 2); op = 2;
            }
            if (h > NEG && h < gbest - XG2) h = NEG;
            H[IX(k)] = h; MM[IX(k)] = mm; GP[IX(k)] = gp; GO[IX(k)] = go; OP[IX(k)] = op;
            if (h > NEG) {
                if (nlo > nhi) nlo = k;
                nhi = k;
                if (h > tmax) tmax = h;
                if (h > best.score2) {
                    best.score2 = h; best.qi = i; best.sj = j;
                    best.mism = mm; best.gaps = gp; best.gapopen = go;
                }
            }
        }
        if (tmax > gbest) gbest = tmax;
        lo2 = lo1; hi2 = hi1; lo1 = nlo; hi1 = nhi;
    }
#undef IX
    free(H); free(MM); free(GP); free(GO); free(OP);
    return best;
}
int64_t orc_dp_cells(void) { return g_dp_cells; }

/* Greedy X-drop extension — the algorithm megablast itself uses for linear gap
 * costs (Zhang, Schwartz, Wagner, Miller 2000; NCBI BLAST_GreedyAlign).  With reward 1, penalty
 * -2 and gap cost 2.5 a path that ends on anti-diagonal t = i + j after d differences
 * (mismatches + gap characters) scores exactly t - 6d in doubled units, so the search keeps, per
 * difference count d and diagonal k = j - i, only the furthest row R(d, k) reachable.
 *   wave 0     : slide from (0, 0) along diagonal 0 while bases match.
 *   wave d >= 1: for every diagonal k in [L-1, U+1] (L..U = valid diagonals of wave d-1) the
 *                entry row is the largest of   R(d-1, k) + 1   mismatch  (needs i<=qn, j<=sn)
 *                                              R(d-1, k+1) + 1 consume q (U; needs i<=qn)
 *                                              R(d-1, k-1)     consume s (L; needs j<=sn)
 *                ties: mismatch, then U, then L.  The entry is dropped when its score
 *                t - 6d < T(d - 19) - 108, T(x) = best score of waves <= x (0 for x < 0):
 *                19 = floor((108 + 1) / 6) + 1 is BLAST's xdrop_offset.  Then slide.
 *   stop       : when a wave has no valid diagonal, or when no diagonal between its lowest and highest valid
 *                one could still beat the best end point even by matching to the end of q or s.
 * Best end point: highest score; ties -> smaller d, then smaller k.  Mismatch / gap / gap-opening
 * counts travel with every (d, k) point: a gap character opens a new gap unless the point it
 * comes from was entered by the same kind of gap and slid over no match. */
static int64_t g_greedy_steps = 0;
static int32_t g_max_band = 0;   /* widest wave (diagonals klo .. khi) of any extension so far: test hook */
int32_t orc_max_band(int reset) { int32_t b = g_max_band; if (reset) g_max_band = 0; return b; }
static ext_t greedy_extend(const uint8_t* q, int32_t qn, int qstep, const uint8_t* s, int32_t sn,
                           int sstep) {
    int32_t nd = qn + sn + 3;
    int32_t* R[2]; int32_t* MM[2]; int32_t* GP[2]; int32_t* GO[2]; uint8_t* OP[2];
    for (int x = 0; x < 2; x++) {
        R[x] = malloc(sizeof(int32_t) * nd); MM[x] = calloc(nd, sizeof(int32_t));
        GP[x] = calloc(nd, sizeof(int32_t)); GO[x] = calloc(nd, sizeof(int32_t)); OP[x] = calloc(nd, 1);
    }
    int32_t tcap = 64, *T = malloc(sizeof(int32_t) * tcap);
#define IX(k) ((k) + qn + 1)
#define QB(i) q[(int64_t)(i) * qstep]
#define SB(j) s[(int64_t)(j) * sstep]
    int32_t i = 0;
    while (i < qn && i < sn && is_match(QB(i), SB(i))) i++;
    g_greedy_steps += i + 1;
    ext_t best = {2 * i, i, i, 0, 0, 0};
    R[0][IX(0)] = i; OP[0][IX(0)] = 0;   /* op: bit0-1 entry op (0 D, 1 U, 2 L), bit2 = slid */
    T[0] = 2 * i;
    int32_t L = 0, U = 0;
    for (int32_t d = 1; d <= qn + sn; d++) {
        const int p = (d - 1) & 1, c = d & 1;
        int32_t tref = d >= 19 ? T[d - 19] : 0;
        int32_t nL = 1, nU = 0, wbest = NEG;
        int32_t klo = L - 1 < -qn ? -qn : L - 1, khi = U + 1 > sn ? sn : U + 1;
        if (khi - klo + 1 > g_max_band) g_max_band = khi - klo + 1;
        for (int32_t k = klo; k <= khi; k++) {
            int32_t bi = -1, mm = 0, gp = 0, go = 0; uint8_t op = 0;
            if (k >= L && k <= U && R[p][IX(k)] >= 0) {
                int32_t ci = R[p][IX(k)] + 1;
                if (ci <= qn && ci + k <= sn) { bi = ci; mm = MM[p][IX(k)] + 1; gp = GP[p][IX(k)]; go = GO[p][IX(k)]; op = 0; }
            }
            if (k + 1 >= L && k + 1 <= U && R[p][IX(k + 1)] >= 0) {
                int32_t ci = R[p][IX(k + 1)] + 1;
                if (ci <= qn && ci > bi) {
                    bi = ci; mm = MM[p][IX(k + 1)]; gp = GP[p][IX(k + 1)] + 1;
                    go = GO[p][IX(k + 1)] + (OP[p][IX(k + 1)] != 1); op = 1;
                }
            }
            if (k - 1 >= L && k - 1 <= U && R[p][IX(k - 1)] >= 0) {
                int32_t ci = R[p][IX(k - 1)];
                if (ci + k <= sn && ci > bi) {
                    bi = ci; mm = MM[p][IX(k - 1)]; gp = GP[p][IX(k - 1)] + 1;
                    go = GO[p][IX(k - 1)] + (OP[p][IX(k - 1)] != 2); op = 2;
                }
            }
            R[c][IX(k)] = -1;
            if (bi < 0) continue;
            if (2 * bi + k - 6 * d < tref - XG2) continue;
            int32_t e = bi;
            while (e < qn && e + k < sn && is_match(QB(e), SB(e + k))) e++;
            g_greedy_steps += e - bi + 1;
            if (e > bi) op |= 4;
            R[c][IX(k)] = e; MM[c][IX(k)] = mm; GP[c][IX(k)] = gp; GO[c][IX(k)] = go; OP[c][IX(k)] = op;
            if (nL > nU) nL = k;
            nU = k;
            int32_t sc = 2 * e + k - 6 * d;
            if (sc > wbest) wbest = sc;
            if (sc > best.score2) { best.score2 = sc; best.qi = e; best.sj = e + k; best.mism = mm; best.gaps = gp; best.gapopen = go; }
        }
        if (d >= tcap) { tcap *= 2; T = realloc(T, sizeof(int32_t) * tcap); }
        T[d] = wbest > T[d - 1] ? wbest : T[d - 1];
        /* wave d-1's cells outside the new range must not be read as wave d+1's predecessors */
        L = nL; U = nU;
        if (L > U) break;
        /* no diagonal of the band L .. U can still beat the best end point (ties go to the smaller d), even by
         * sliding to the end of q or s without a further difference: 2 min(qn, sn - k) + k - 6 d peaks at
         * k = sn - qn.  The result is final; this saves the waves that only run down the X-drop after a
         * full-length alignment has ended. */
        {
            int32_t kp = sn - qn < L ? L : (sn - qn > U ? U : sn - qn);
            int32_t rowmax = qn < sn - kp ? qn : sn - kp;
            if (2 * rowmax + kp - 6 * d <= best.score2) break;
        }
    }
#undef IX
#undef QB
#undef SB
    for (int x = 0; x < 2; x++) { free(R[x]); free(MM[x]); free(GP[x]); free(GO[x]); free(OP[x]); }
    free(T);
    return best;
}
int64_t orc_greedy_steps(void) { return g_greedy_steps; }
static int g_gapped_mode = 1;   /* 1 = greedy (canonical), 0 = anti-diagonal X-drop DP (cross-check) */
void orc_set_gapped_mode(int m) { g_gapped_mode = m; }
static ext_t gapped_extend(const uint8_t* q, int32_t qn, int qstep, const uint8_t* s, int32_t sn, int sstep) {
    return g_gapped_mode ? greedy_extend(q, qn, qstep, s, sn, sstep) : xdrop_extend(q, qn, qstep, s, sn, sstep);
}

/* test hook: one extension of ASCII q against ASCII s (both forward); out = score2 qi sj mism gaps gapopen */
void orc_extend(const uint8_t* q, int32_t qn, const uint8_t* s, int32_t sn, int mode, int32_t* out) {
    uint8_t* a = malloc(qn + 1); uint8_t* b = malloc(sn + 1);
    for (int32_t i = 0; i < qn; i++) a[i] = code_of(q[i]);
    for (int32_t i = 0; i < sn; i++) b[i] = code_of(s[i]);
    ext_t r = mode ? greedy_extend(a, qn, 1, b, sn, 1) : xdrop_extend(a, qn, 1, b, sn, 1);
    out[0] = r.score2; out[1] = r.qi; out[2] = r.sj; out[3] = r.mism; out[4] = r.gaps; out[5] = r.gapopen;
    free(a); free(b);
}

/* ------------------------------- seeds -> ungapped HSPs ------------------------------- */
typedef struct { uint64_t key; int32_t tb, pos; int32_t minus; } kent_t;
static int cmp_kent(const void* x, const void* y) {
    const kent_t *a = x, *b = y;
    if (a->key != b->key) return a->key < b->key ? -1 : 1;
    if (a->minus != b->minus) return a->minus - b->minus;
    if (a->tb != b->tb) return a->tb - b->tb;
    return a->pos - b->pos;
}
static int cmp_ungapped_id(const void* x, const void* y) {
    const ungapped_t *a = x, *b = y;
    if (a->ta != b->ta) return a->ta < b->ta ? -1 : 1;
    if (a->tb != b->tb) return a->tb < b->tb ? -1 : 1;
    if (a->minus != b->minus) return a->minus - b->minus;
    if (a->diag != b->diag) return a->diag < b->diag ? -1 : 1;
    if (a->a0 != b->a0) return a->a0 < b->a0 ? -1 : 1;
    if (a->a1 != b->a1) return a->a1 < b->a1 ? -1 : 1;
    return 0;
}
/* processing order inside a group: score desc, diag asc, a0 asc, a1 asc */
static int cmp_ungapped_order(const void* x, const void* y) {
    const ungapped_t *a = x, *b = y;
    if (a->ta != b->ta) return a->ta < b->ta ? -1 : 1;
    if (a->tb != b->tb) return a->tb < b->tb ? -1 : 1;
    if (a->minus != b->minus) return a->minus - b->minus;
    if (a->score != b->score) return a->score > b->score ? -1 : 1;
    if (a->diag != b->diag) return a->diag < b->diag ? -1 : 1;
    if (a->a0 != b->a0) return a->a0 < b->a0 ? -1 : 1;
    if (a->a1 != b->a1) return a->a1 < b->a1 ? -1 : 1;
    return 0;
}

/* rolling 28-mer keys of one sequence; valid[i] = window [i, i+28) has no N and no DUST-masked base */
static void kmer_keys(const uint8_t* s, const uint8_t* dust, int32_t n, uint64_t* key, uint8_t* valid) {
    uint64_t k = 0; int32_t good = 0;
    const uint64_t maskk = (1ull << (2 * WORD)) - 1;
    for (int32_t i = 0; i < n; i++) {
        if (s[i] > 3) { good = 0; k = 0; } else { good = dust[i] ? 0 : good + 1; k = ((k << 2) | s[i]) & maskk; }
        if (i >= WORD - 1) { key[i - WORD + 1] = k; valid[i - WORD + 1] = good >= WORD; }
    }
}

typedef struct { ungapped_t* v; int64_t n, cap; } uvec_t;
static void upush(uvec_t* u, ungapped_t x) {
    if (u->n == u->cap) { u->cap = u->cap ? u->cap * 2 : 1024; u->v = realloc(u->v, u->cap * sizeof(ungapped_t)); }
    u->v[u->n++] = x;
}

static int64_t g_seeds = 0;
/* Seeds: all maximal exact runs >= WORD between transcripts of s1 (plus) and s2 (both strands) that hold
 * at least one window of WORD bases free of DUST-masked bases in BOTH samples (blastn masks the query of
 * each direction; one search serves both directions here, so both masks apply).  Each seed is then
 * extended without gaps under BLAST's diagonal rule (below); duplicates removed. */
static uvec_t find_ungapped(const orc_sample* s1, const orc_sample* s2) {
    int64_t cap = 2 * s2->total + 1, nk = 0;
    kent_t* idx = malloc(cap * sizeof(kent_t));
    int32_t maxlen = 1;
    for (int t = 0; t < s1->n_seq; t++) if (s1->off[t + 1] - s1->off[t] > maxlen) maxlen = (int32_t)(s1->off[t + 1] - s1->off[t]);
    for (int t = 0; t < s2->n_seq; t++) if (s2->off[t + 1] - s2->off[t] > maxlen) maxlen = (int32_t)(s2->off[t + 1] - s2->off[t]);
    uint64_t* key = malloc(sizeof(uint64_t) * (maxlen + 1));
    uint8_t* valid = malloc(maxlen + 1);
    for (int t = 0; t < s2->n_seq; t++) {
        int32_t n = (int32_t)(s2->off[t + 1] - s2->off[t]);
        for (int minus = 0; minus < 2; minus++) {
            const uint8_t* b = (minus ? s2->rc : s2->fwd) + s2->off[t];
            if (n < WORD) continue;
            kmer_keys(b, (minus ? s2->dust_rc : s2->dust) + s2->off[t], n, key, valid);
            for (int32_t j = 0; j + WORD <= n; j++)
                if (valid[j]) { idx[nk].key = key[j]; idx[nk].tb = t; idx[nk].pos = j; idx[nk].minus = minus; nk++; }
        }
    }
    qsort(idx, nk, sizeof(kent_t), cmp_kent);
    /* pass 1: every seed = maximal exact run holding a clean window, as (ta, tb, minus, diag, [rs, re)):
     * each stretch of consecutive clean matching windows reports its run once; a run with several
     * stretches is reported several times and collapses in the sort below */
    uvec_t seeds = {0, 0, 0};
    for (int ta = 0; ta < s1->n_seq; ta++) {
        const uint8_t* a = s1->fwd + s1->off[ta];
        const uint8_t* da = s1->dust + s1->off[ta];
        int32_t la = (int32_t)(s1->off[ta + 1] - s1->off[ta]);
        if (la < WORD) continue;
        kmer_keys(a, da, la, key, valid);
        for (int32_t i = 0; i + WORD <= la; i++) {
            if (!valid[i]) continue;
            int64_t lo = 0, hi = nk;
            while (lo < hi) { int64_t mid = (lo + hi) / 2; if (idx[mid].key < key[i]) lo = mid + 1; else hi = mid; }
            for (int64_t e = lo; e < nk && idx[e].key == key[i]; e++) {
                int32_t tb = idx[e].tb, j = idx[e].pos, minus = idx[e].minus;
                const uint8_t* b = (minus ? s2->rc : s2->fwd) + s2->off[tb];
                const uint8_t* db = (minus ? s2->dust_rc : s2->dust) + s2->off[tb];
                int32_t lb = (int32_t)(s2->off[tb + 1] - s2->off[tb]);
                /* the window one base to the left also matches and is clean: not the first of its stretch */
                if (i > 0 && j > 0 && is_match(a[i - 1], b[j - 1]) && !da[i - 1] && !db[j - 1]) continue;
                int32_t rs = i, re = i + WORD;
                while (rs > 0 && rs + (j - i) > 0 && is_match(a[rs - 1], b[rs - 1 + (j - i)])) rs--;
                while (re < la && re + (j - i) < lb && is_match(a[re], b[re + (j - i)])) re++;
                ungapped_t u;
                u.ta = ta; u.tb = tb; u.minus = minus; u.diag = j - i; u.a0 = rs; u.a1 = re; u.score = 0;
                upush(&seeds, u);
            }
        }
    }
    free(idx); free(key); free(valid);
    /* pass 2 (BLAST's diagonal rule): along each diagonal, seeds in ascending position; a seed that
     * starts inside the ungapped HSP extended from an earlier seed of the diagonal is skipped */
    uvec_t out = {0, 0, 0};
    if (seeds.n) {
        qsort(seeds.v, seeds.n, sizeof(ungapped_t), cmp_ungapped_id);
        int64_t w = 1;
        for (int64_t r = 1; r < seeds.n; r++)
            if (cmp_ungapped_id(&seeds.v[r], &seeds.v[w - 1]) != 0) seeds.v[w++] = seeds.v[r];
        seeds.n = w;
    }
    g_seeds += seeds.n;
    int32_t last_end = -1;
    for (int64_t r = 0; r < seeds.n; r++) {
        const ungapped_t* sd = &seeds.v[r];
        if (r == 0 || sd->ta != seeds.v[r - 1].ta || sd->tb != seeds.v[r - 1].tb ||
            sd->minus != seeds.v[r - 1].minus || sd->diag != seeds.v[r - 1].diag) last_end = -1;
        if (sd->a0 < last_end) continue;
        const uint8_t* a = s1->fwd + s1->off[sd->ta];
        const uint8_t* b = (sd->minus ? s2->rc : s2->fwd) + s2->off[sd->tb];
        int32_t la = (int32_t)(s1->off[sd->ta + 1] - s1->off[sd->ta]);
        int32_t lb = (int32_t)(s2->off[sd->tb + 1] - s2->off[sd->tb]);
        ungapped_t u = ungapped_extend(a, la, b, lb, sd->a0, sd->a1, sd->diag);
        u.ta = sd->ta; u.tb = sd->tb; u.minus = sd->minus;
        last_end = u.a1;
        upush(&out, u);
    }
    free(seeds.v);
    if (out.n) {
        qsort(out.v, out.n, sizeof(ungapped_t), cmp_ungapped_id);
        int64_t w = 1;
        for (int64_t r = 1; r < out.n; r++)
            if (cmp_ungapped_id(&out.v[r], &out.v[w - 1]) != 0) out.v[w++] = out.v[r];
        out.n = w;
        qsort(out.v, out.n, sizeof(ungapped_t), cmp_ungapped_order);
    }
    return out;
}

/* ------------------------------- gapped stage ----------------------------------------- */
typedef struct { gapped_t* v; int64_t n, cap; } gvec_t;
static void gpush(gvec_t* g, gapped_t x) {
    if (g->n == g->cap) { g->cap = g->cap ? g->cap * 2 : 1024; g->v = realloc(g->v, g->cap * sizeof(gapped_t)); }
    g->v[g->n++] = x;
}

static gapped_t gapped_from(const orc_sample* s1, const orc_sample* s2, const ungapped_t* u) {
    const uint8_t* a = s1->fwd + s1->off[u->ta];
    const uint8_t* b = (u->minus ? s2->rc : s2->fwd) + s2->off[u->tb];
    int32_t la = (int32_t)(s1->off[u->ta + 1] - s1->off[u->ta]);
    int32_t lb = (int32_t)(s2->off[u->tb + 1] - s2->off[u->tb]);
    int32_t i0 = gapped_start(a, b, u), j0 = i0 + u->diag;
    /* left part consumes the anchor pair (i0, j0) first, walking backwards */
    ext_t L = gapped_extend(a + i0, i0 + 1, -1, b + j0, j0 + 1, -1);
    ext_t R = gapped_extend(a + i0 + 1, la - i0 - 1, 1, b + j0 + 1, lb - j0 - 1, 1);
    gapped_t g;
    g.ta = u->ta; g.tb = u->tb; g.minus = u->minus;
    g.a0 = i0 + 1 - L.qi; g.a1 = i0 + 1 + R.qi;
    g.b0 = j0 + 1 - L.sj; g.b1 = j0 + 1 + R.sj;
    g.score2 = L.score2 + R.score2;
    g.mism = L.mism + R.mism; g.gaps = L.gaps + R.gaps; g.gapopen = L.gapopen + R.gapopen;
    g.nident = ((g.a1 - g.a0) + (g.b1 - g.b0) - g.gaps) / 2 - g.mism;
    return g;
}

static gvec_t find_gapped(const orc_sample* s1, const orc_sample* s2, const uvec_t* U) {
    gvec_t G = {0, 0, 0};
    int64_t r = 0;
    while (r < U->n) {
        int64_t e = r;
        while (e < U->n && U->v[e].ta == U->v[r].ta && U->v[e].tb == U->v[r].tb && U->v[e].minus == U->v[r].minus) e++;
        int64_t g0 = G.n;
        for (int64_t x = r; x < e && G.n - g0 < MAXG; x++) {
            const ungapped_t* u = &U->v[x];
            int32_t ub0 = u->a0 + u->diag, ub1 = u->a1 + u->diag;
            int skip = 0;
            for (int64_t y = g0; y < G.n && !skip; y++) {
                const gapped_t* g = &G.v[y];
                if (g->a0 <= u->a0 && u->a1 <= g->a1 && g->b0 <= ub0 && ub1 <= g->b1) skip = 1;
            }
            if (skip) continue;
            gapped_t g = gapped_from(s1, s2, u);
            for (int64_t y = g0; y < G.n && !skip; y++) {
                const gapped_t* h = &G.v[y];
                if (h->a0 == g.a0 && h->a1 == g.a1 && h->b0 == g.b0 && h->b1 == g.b1) skip = 1;
            }
            if (!skip) gpush(&G, g);
        }
        r = e;
    }
    return G;
}

/* ------------------------------- table assembly --------------------------------------- */
typedef struct {
    int32_t length, mismatch, gapopen, qstart, qend, sstart, send, score, gaps, nident;
    int32_t qgene, qiso, sgene, siso, qlen;
    int32_t minus, reverse;
    int32_t pos;  /* position in the frame the row currently lives in (stable tie order) */
} row_t;
typedef struct { row_t* v; int64_t n, cap; } rvec_t;
static void rpush(rvec_t* r, row_t x) {
    if (r->n == r->cap) { r->cap = r->cap ? r->cap * 2 : 1024; r->v = realloc(r->v, r->cap * sizeof(row_t)); }
    r->v[r->n++] = x;
}

static int passes_evalue(int32_t score, int64_t m, int64_t n, int64_t nseq, double log_cut) {
    int32_t la = orc_length_adjustment(m, n, nseq);
    double em = (double)(m - la), en = (double)(n - nseq * (int64_t)la);
    if (em < 1) em = 1;
    if (en < 1) en = 1;
    return log(KA_K * em * en) - KA_LAMBDA * (double)score <= log_cut;
}

/* One blastn direction as rows.  forward: query = sample 2, subject = sample 1
 * (find_homologs.py:229-232 with :124 TabularBlastnSearch(path2, path1): the table contract of
 * docs/formats.md:246-252 — qgene is a gene of the second sample).  backward: roles swapped. */
static rvec_t direction_rows(const orc_sample* s1, const orc_sample* s2, const gvec_t* G,
                             int backward, double log_cut) {
    rvec_t R = {0, 0, 0};
    for (int64_t x = 0; x < G->n; x++) {
        const gapped_t* g = &G->v[x];
        int32_t la = (int32_t)(s1->off[g->ta + 1] - s1->off[g->ta]);
        int32_t lb = (int32_t)(s2->off[g->tb + 1] - s2->off[g->tb]);
        (void)la;
        row_t r; memset(&r, 0, sizeof r);
        r.score = g->score2 >> 1;
        r.gaps = g->gaps; r.mismatch = g->mism; r.gapopen = g->gapopen; r.nident = g->nident;
        r.length = g->nident + g->mism + g->gaps;
        r.minus = g->minus; r.reverse = backward;
        /* plus-strand coordinates of the sample-2 range */
        int32_t p0 = g->minus ? lb - g->b1 : g->b0, p1 = g->minus ? lb - g->b0 : g->b1;
        if (!backward) {
            r.qgene = s2->gene[g->tb]; r.qiso = s2->iso[g->tb];
            r.sgene = s1->gene[g->ta]; r.siso = s1->iso[g->ta];
            r.qlen = lb;
            r.qstart = p0 + 1; r.qend = p1;
            if (!g->minus) { r.sstart = g->a0 + 1; r.send = g->a1; }
            else { r.sstart = g->a1; r.send = g->a0 + 1; }
            if (!passes_evalue(r.score, lb, s1->total, s1->n_seq, log_cut)) continue;
        } else {
            r.qgene = s1->gene[g->ta]; r.qiso = s1->iso[g->ta];
            r.sgene = s2->gene[g->tb]; r.siso = s2->iso[g->tb];
            r.qlen = (int32_t)(s1->off[g->ta + 1] - s1->off[g->ta]);
            r.qstart = g->a0 + 1; r.qend = g->a1;
            if (!g->minus) { r.sstart = p0 + 1; r.send = p1; }
            else { r.sstart = p1; r.send = p0 + 1; }
            if (!passes_evalue(r.score, r.qlen, s2->total, s2->n_seq, log_cut)) continue;
        }
        rpush(&R, r);
    }
    return R;
}

/* highest_bitscores (find_homologs.py:135-164): per group the rows holding the n largest
 * scores; keep_all -> every row tied with the n-th is kept (pandas nlargest keep="all"),
 * else exactly the first n in (score desc, current position asc) order (keep="first").
 * Output order = pandas': group key ascending, then score descending, then position.
 * group: 0 = qgene, 1 = (qgene, sgene). */
static int g_group_mode = 0;
static int cmp_rows_group(const void* x, const void* y) {
    const row_t *a = x, *b = y;
    if (a->qgene != b->qgene) return a->qgene < b->qgene ? -1 : 1;
    if (g_group_mode == 1 && a->sgene != b->sgene) return a->sgene < b->sgene ? -1 : 1;
    if (a->score != b->score) return a->score > b->score ? -1 : 1;
    if (a->pos != b->pos) return a->pos < b->pos ? -1 : 1;
    return 0;
}
static void select_topn(rvec_t* R, int group_mode, int n, int keep_all) {
    for (int64_t i = 0; i < R->n; i++) R->v[i].pos = (int32_t)i;
    g_group_mode = group_mode;
    qsort(R->v, R->n, sizeof(row_t), cmp_rows_group);
    int64_t w = 0, r = 0;
    while (r < R->n) {
        int64_t e = r;
        while (e < R->n && R->v[e].qgene == R->v[r].qgene && (group_mode == 0 || R->v[e].sgene == R->v[r].sgene)) e++;
        int64_t take = (e - r < n) ? e - r : n;
        if (keep_all) { int32_t thr = R->v[r + take - 1].score; while (r + take < e && R->v[r + take].score == thr) take++; }
        for (int64_t k = 0; k < take; k++) R->v[w++] = R->v[r + k];
        r = e;
    }
    R->n = w;
}

typedef struct { int32_t q, s; } pair_t;
static int cmp_pair(const void* x, const void* y) {
    const pair_t *a = x, *b = y;
    if (a->q != b->q) return a->q < b->q ? -1 : 1;
    if (a->s != b->s) return a->s < b->s ? -1 : 1;
    return 0;
}
static int has_pair(const pair_t* p, int64_t n, int32_t q, int32_t s) {
    pair_t k = {q, s};
    return bsearch(&k, p, n, sizeof(pair_t), cmp_pair) != NULL;
}

/* HomologFinder.get_match_table (find_homologs.py:215-302) on two direction frames that have
 * already been reduced by select_topn(group=qgene, keep_all=1).  The backward frame is renamed
 * (q <-> s for gene/iso) first, so afterwards qgene is always a gene of sample 2. */
static rvec_t match_table(rvec_t* F, rvec_t* B, int n, int keep_all) {
    for (int64_t i = 0; i < B->n; i++) {
        row_t* r = &B->v[i];
        int32_t t = r->qgene; r->qgene = r->sgene; r->sgene = t;
        t = r->qiso; r->qiso = r->siso; r->siso = t;
    }
    pair_t* pf = malloc(sizeof(pair_t) * (F->n + 1));
    pair_t* pb = malloc(sizeof(pair_t) * (B->n + 1));
    for (int64_t i = 0; i < F->n; i++) { pf[i].q = F->v[i].qgene; pf[i].s = F->v[i].sgene; }
    for (int64_t i = 0; i < B->n; i++) { pb[i].q = B->v[i].qgene; pb[i].s = B->v[i].sgene; }
    qsort(pf, F->n, sizeof(pair_t), cmp_pair);
    qsort(pb, B->n, sizeof(pair_t), cmp_pair);
    rvec_t T = {0, 0, 0};
    for (int64_t i = 0; i < F->n; i++) if (has_pair(pb, B->n, F->v[i].qgene, F->v[i].sgene)) rpush(&T, F->v[i]);
    for (int64_t i = 0; i < B->n; i++) if (has_pair(pf, F->n, B->v[i].qgene, B->v[i].sgene)) rpush(&T, B->v[i]);
    free(pf); free(pb);
    /* the reference calls highest_bitscores here WITHOUT n (find_homologs.py:281-301), so both
     * steps use n = 1 whatever top_n is; top_n only shapes the two direction frames */
    (void)n;
    select_topn(&T, 1, 1, 1);
    select_topn(&T, 0, 1, keep_all);
    return T;
}

/* ------------------------------- exported entry points -------------------------------- */
typedef struct {
    gvec_t G;        /* symmetric gapped HSP set */
    rvec_t F, B, T;  /* forward / backward direction frames after per-gene selection; table */
    int64_t n_seeds, n_ungapped, dp_cells;
} orc_result;

orc_result* orc_search_pair(const orc_sample* s1, const orc_sample* s2, double evalue, int top_n,
                            int keep_all) {
    orc_result* res = calloc(1, sizeof *res);
    int64_t c0 = g_dp_cells, sd0 = g_seeds, gs0 = g_greedy_steps;
    uvec_t U = find_ungapped(s1, s2);
    res->n_ungapped = U.n;
    res->G = find_gapped(s1, s2, &U);
    free(U.v);
    double log_cut = evalue > 0 ? log(evalue) : -INFINITY;
    res->F = direction_rows(s1, s2, &res->G, 0, log_cut);
    res->B = direction_rows(s1, s2, &res->G, 1, log_cut);
    select_topn(&res->F, 0, top_n, 1);
    select_topn(&res->B, 0, top_n, 1);
    rvec_t Fc = res->F, Bc = res->B;
    Fc.v = malloc(sizeof(row_t) * (Fc.n + 1)); memcpy(Fc.v, res->F.v, sizeof(row_t) * Fc.n);
    Bc.v = malloc(sizeof(row_t) * (Bc.n + 1)); memcpy(Bc.v, res->B.v, sizeof(row_t) * Bc.n);
    res->T = match_table(&Fc, &Bc, top_n, keep_all);
    free(Fc.v); free(Bc.v);
    res->dp_cells = g_gapped_mode ? g_greedy_steps - gs0 : g_dp_cells - c0;
    res->n_seeds = g_seeds - sd0;
    return res;
}
/* The two blastn runs of find_homologs.py:229-239 as two INDEPENDENT searches, each with its own query
 * role: forward = query sample 2 against database sample 1, backward = query sample 1 against
 * database sample 2.  In each the query transcript plays the part of `q` in the extensions (gap tie
 * rule, start point, diagonal rule run in the query's own orientation), exactly as in the backward
 * search of orc_search_pair.  orc_search_pair derives BOTH frames from the one search with sample 1 as
 * the query (the product's shortcut: one search per unordered pair); this entry measures what the
 * shortcut changes (tools/direction_asymmetry.py, DESIGN.md section 3 "Not modelled"). */
orc_result* orc_search_pair_directions(const orc_sample* s1, const orc_sample* s2, double evalue, int top_n,
                                       int keep_all) {
    orc_result* res = calloc(1, sizeof *res);
    int64_t c0 = g_dp_cells, sd0 = g_seeds, gs0 = g_greedy_steps;
    double log_cut = evalue > 0 ? log(evalue) : -INFINITY;
    uvec_t U = find_ungapped(s1, s2);
    res->n_ungapped = U.n;
    res->G = find_gapped(s1, s2, &U);          /* backward search: query = sample 1 */
    free(U.v);
    res->B = direction_rows(s1, s2, &res->G, 1, log_cut);
    uvec_t U2 = find_ungapped(s2, s1);         /* forward search: query = sample 2 */
    gvec_t Gf = find_gapped(s2, s1, &U2);
    free(U2.v);
    res->F = direction_rows(s2, s1, &Gf, 1, log_cut);   /* rows with q = sample 2, s = sample 1 */
    for (int64_t i = 0; i < res->F.n; i++) res->F.v[i].reverse = 0;
    free(Gf.v);
    select_topn(&res->F, 0, top_n, 1);
    select_topn(&res->B, 0, top_n, 1);
    rvec_t Fc = res->F, Bc = res->B;
    Fc.v = malloc(sizeof(row_t) * (Fc.n + 1)); memcpy(Fc.v, res->F.v, sizeof(row_t) * Fc.n);
    Bc.v = malloc(sizeof(row_t) * (Bc.n + 1)); memcpy(Bc.v, res->B.v, sizeof(row_t) * Bc.n);
    res->T = match_table(&Fc, &Bc, top_n, keep_all);
    free(Fc.v); free(Bc.v);
    res->dp_cells = g_gapped_mode ? g_greedy_steps - gs0 : g_dp_cells - c0;
    res->n_seeds = g_seeds - sd0;
    return res;
}
void orc_result_free(orc_result* r) {
    if (!r) return;
    free(r->G.v); free(r->F.v); free(r->B.v); free(r->T.v); free(r);
}
/* counts: [0]=#gapped HSPs [1]=#fwd rows [2]=#bwd rows [3]=#table rows [4]=#seeds [5]=#ungapped
 * [6]=base comparisons of the gapped stage ("extension steps": the greedy waves' slide steps; cells of the
 * cross-check DP when that mode is selected) */
void orc_result_counts(const orc_result* r, int64_t* c) {
    c[0] = r->G.n; c[1] = r->F.n; c[2] = r->B.n; c[3] = r->T.n;
    c[4] = r->n_seeds; c[5] = r->n_ungapped; c[6] = r->dp_cells;
}
/* 12 int32 per HSP, same layout as rcq_debug_last_hsps (include/rcq.h) */
void orc_result_hsps(const orc_result* r, int32_t* out) {
    for (int64_t i = 0; i < r->G.n; i++) {
        const gapped_t* g = &r->G.v[i];
        int32_t* o = out + 12 * i;
        o[0] = g->ta; o[1] = g->tb; o[2] = g->minus; o[3] = g->a0; o[4] = g->a1; o[5] = g->b0;
        o[6] = g->b1; o[7] = g->score2; o[8] = g->mism; o[9] = g->gaps; o[10] = g->gapopen; o[11] = g->nident;
    }
}
/* which: 0 = forward frame, 1 = backward frame, 2 = table.  17 int32 per row:
 * length mismatch gapopen qstart qend sstart send score gaps nident qgene qiso sgene siso qlen minus reverse */
void orc_result_rows(const orc_result* r, int which, int32_t* out) {
    const rvec_t* R = which == 0 ? &r->F : which == 1 ? &r->B : &r->T;
    for (int64_t i = 0; i < R->n; i++) {
        const row_t* x = &R->v[i];
        int32_t* o = out + 17 * i;
        o[0] = x->length; o[1] = x->mismatch; o[2] = x->gapopen; o[3] = x->qstart; o[4] = x->qend;
        o[5] = x->sstart; o[6] = x->send; o[7] = x->score; o[8] = x->gaps; o[9] = x->nident;
        o[10] = x->qgene; o[11] = x->qiso; o[12] = x->sgene; o[13] = x->siso; o[14] = x->qlen;
        o[15] = x->minus; o[16] = x->reverse;
    }
}
/* Table assembly alone, from caller-supplied direction frames (17 int32 per row as above, BEFORE
 * the per-gene selection) — the entry the golden test drives with the same hit frames it gave
 * the reference's HomologFinder.get_match_table. */
orc_result* orc_match_table_from_rows(const int32_t* fwd, int64_t nf, const int32_t* bwd, int64_t nb,
                                      int top_n, int keep_all) {
    orc_result* res = calloc(1, sizeof *res);
    const int32_t* src[2] = {fwd, bwd};
    int64_t cnt[2] = {nf, nb};
    rvec_t* dst[2] = {&res->F, &res->B};
    for (int d = 0; d < 2; d++)
        for (int64_t i = 0; i < cnt[d]; i++) {
            const int32_t* o = src[d] + 17 * i;
            row_t x; memset(&x, 0, sizeof x);
            x.length = o[0]; x.mismatch = o[1]; x.gapopen = o[2]; x.qstart = o[3]; x.qend = o[4];
            x.sstart = o[5]; x.send = o[6]; x.score = o[7]; x.gaps = o[8]; x.nident = o[9];
            x.qgene = o[10]; x.qiso = o[11]; x.sgene = o[12]; x.siso = o[13]; x.qlen = o[14];
            x.minus = o[15]; x.reverse = d;
            rpush(dst[d], x);
        }
    select_topn(&res->F, 0, top_n, 1);
    select_topn(&res->B, 0, top_n, 1);
    rvec_t Fc = res->F, Bc = res->B;
    Fc.v = malloc(sizeof(row_t) * (Fc.n + 1)); memcpy(Fc.v, res->F.v, sizeof(row_t) * Fc.n);
    Bc.v = malloc(sizeof(row_t) * (Bc.n + 1)); memcpy(Bc.v, res->B.v, sizeof(row_t) * Bc.n);
    res->T = match_table(&Fc, &Bc, top_n, keep_all);
    free(Fc.v); free(Bc.v);
    return res;
}
