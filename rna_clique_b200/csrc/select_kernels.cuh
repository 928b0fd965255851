// select_kernels.cuh — kernel (3c): e-value cut, per-query-gene best hit of both blastn directions and
// HomologFinder.get_match_table (reference find_homologs.py:135-164, 215-302) on the device.
//
// Input: the symmetric gapped-HSP list G of one sample pair (A = sample 1 = transcripts1 =
// "ssample" side, B = sample 2 = "qsample" side).  forward search: query B, subject A;
// backward search: query A, subject B.  After the backward frame's q/s rename every row has
// qgene in B and sgene in A.
//   F  = forward rows holding the top_n largest scores of their query gene, every row tied with the
//        n-th kept (pandas nlargest(n, keep="all"), find_homologs.py:135-164); top_n = 1 (the reference
//        default, rna_clique.py:37) is a per-gene atomicMax, top_n > 1 a sort by (gene, score desc) whose
//        n-th row of every gene gives that gene's threshold
//   Bk = the same for the backward rows, grouped by THEIR query gene (a gene of A)
//   keep rows whose (qgene, sgene) pair occurs in both; then per pair the best score (ties kept),
//   then per qgene the best score (keep_all) or the first such row (keep="first").
// An HSP that qualifies in both directions yields two rows (reverse = 0 and 1) — the duplicate
// rows the reference produces when forward and backward bitscores tie.
#pragma once
#include "search.cuh"

#define EMPTY64 0xFFFFFFFFFFFFFFFFull

// BLAST_ComputeLengthAdjustment for blastn 1/-2 linear gaps (same arithmetic as the oracle)
__host__ __device__ inline int32_t length_adjustment(double m, double n, double N) {
    const double K = RCQ_KA_K, logK = log(RCQ_KA_K), adl = RCQ_KA_ALPHA / RCQ_KA_LAMBDA, beta = RCQ_KA_BETA;
    double ell, ss, ell_bar, ell_min = 0, ell_max, ell_next = 0;
    bool converged = false;
    double a = N, mb = m * N + n, c = n * m - (m > n ? m : n) / K;
    if (c < 0) return 0;
    ell_max = 2 * c / (mb + sqrt(mb * mb - 4 * a * c));
    for (int i = 1; i <= 20; i++) {
        ell = ell_next;
        ss = (m - ell) * (n - N * ell);
        ell_bar = adl * (logK + log(ss)) + beta;
        if (ell_bar >= ell) {
            ell_min = ell;
            if (ell_bar - ell_min <= 1.0) { converged = true; break; }
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
__host__ __device__ inline double log_evalue(int32_t score, int64_t m, int64_t n, int64_t nseq) {
    int32_t la = length_adjustment((double)m, (double)n, (double)nseq);
    double em = (double)(m - la), en = (double)(n - nseq * (int64_t)la);
    if (em < 1) em = 1;
    if (en < 1) en = 1;
    return log(RCQ_KA_K * em * en) - RCQ_KA_LAMBDA * (double)score;
}

extern "C" int32_t rcq_length_adjustment(int64_t m, int64_t n, int64_t nseq) {
    return length_adjustment((double)m, (double)n, (double)nseq);
}
extern "C" double rcq_evalue(int32_t score, int64_t m, int64_t n, int64_t nseq) {
    return exp(log_evalue(score, m, n, nseq));
}
extern "C" double rcq_bitscore(int32_t score) { return (RCQ_KA_LAMBDA * score - log(RCQ_KA_K)) / log(2.0); }

// The three floating-point columns of a gene-matches table as blastn prints them and simple_blast parses them
// back (find_homologs.py:124-125 reads them from the tabular output): pident = 100 nident / length to three
// decimals; evalue = K m' n' exp(-lambda S) with the length adjustment of the row's direction, printed as 0.0
// below 1e-180, with three significant digits below 0.0009, in BLAST's fixed formats above; bitscore truncated to
// an integer above 99.9 bits, one decimal below.  Host code, one pass; the same arithmetic, in the same order, as
// find_homologs.table_stats_numpy (tests compare the two).  ln_k, ln_2: the caller's log(K), log(2).
extern "C" int rcq_table_stats(int64_t n, const int32_t* score, const int32_t* qlen, const uint8_t* reverse,
                               const int32_t* nident, const int32_t* length, int64_t bases_fwd, int64_t seqs_fwd,
                               int64_t bases_bwd, int64_t seqs_bwd, double ln_k, double ln_2, double* pident,
                               double* evalue, double* bitscore) {
    if (n < 0 || (n && (!score || !qlen || !reverse || !nident || !length || !pident || !evalue || !bitscore)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_table_stats: bad argument");
    // the length adjustment depends on (query length, direction) only: a direct-mapped cache by length
    constexpr int CACHE = 4096;
    std::vector<int32_t> ckey(2 * CACHE, -1);
    std::vector<double> cla(2 * CACHE, 0.0);
    for (int64_t i = 0; i < n; i++) {
        const int r = reverse[i] ? 1 : 0;
        const int slot = r * CACHE + (qlen[i] & (CACHE - 1));
        if (ckey[slot] != qlen[i]) {
            ckey[slot] = qlen[i];
            cla[slot] = (double)length_adjustment((double)qlen[i], (double)(r ? bases_bwd : bases_fwd), (double)(r ? seqs_bwd : seqs_fwd));
        }
        const double la = cla[slot];
        const double nb = (double)(r ? bases_bwd : bases_fwd), ns = (double)(r ? seqs_bwd : seqs_fwd);
        double em = (double)qlen[i] - la, en = nb - ns * la;
        if (em < 1.0) em = 1.0;
        if (en < 1.0) en = 1.0;
        const double e = RCQ_KA_K * em * en * exp(-RCQ_KA_LAMBDA * (double)score[i]);
        double pe;
        if (e < 1e-180) pe = 0.0;
        else if (e < 0.0009) { char buf[40]; snprintf(buf, sizeof buf, "%.2e", e); pe = strtod(buf, nullptr); }
        else if (e < 0.1) pe = nearbyint(e * 1000.0) / 1000.0;
        else if (e < 1.0) pe = nearbyint(e * 100.0) / 100.0;
        else if (e < 10.0) pe = nearbyint(e * 10.0) / 10.0;
        else pe = nearbyint(e);
        evalue[i] = pe;
        const double b = (RCQ_KA_LAMBDA * (double)score[i] - ln_k) / ln_2;
        bitscore[i] = b > 99.9 ? floor(b) : nearbyint(b * 10.0) / 10.0;
        const int32_t len = length[i] > 1 ? length[i] : 1;
        pident[i] = nearbyint(100.0 * (double)nident[i] / (double)len * 1000.0) / 1000.0;
    }
    return 0;
}

struct SelView {
    const Gapped* G; int64_t n;
    const int32_t *lenA, *lenB, *gidxA, *gidxB;
    int64_t totalA, totalB; int32_t nA, nB;
    double log_cut;
    int32_t top_n;
    uint8_t* pass;       // bit0 forward passes e-value, bit1 backward
    int32_t *bestF, *bestB, *bestQ;   // per gene of B / A / B; bestF/bestB = lowest score still inside the gene's top_n
    uint64_t* hkeys; uint32_t* hflags; int32_t* hbest; uint32_t hmask;  // (qgene,sgene) pair set
    unsigned long long* firstkey;     // per gene of B (keep="first")
};

__device__ __forceinline__ uint32_t pair_slot(const SelView& V, uint64_t key, bool insert) {
    uint32_t s = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32) & V.hmask;
    while (true) {
        uint64_t cur = V.hkeys[s];
        if (cur == key) return s;
        if (cur == EMPTY64) {
            if (!insert) return 0xFFFFFFFFu;
            uint64_t old = atomicCAS((unsigned long long*)&V.hkeys[s], EMPTY64, key);
            if (old == EMPTY64 || old == key) return s;
        }
        s = (s + 1) & V.hmask;
    }
}

// pass 1: e-value flags and per-query-gene best score in each direction
__global__ void k_sel_best(SelView V) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    int32_t raw = g.score2 >> 1;
    int32_t la = V.lenA[g.ta], lb = V.lenB[g.tb];
    uint8_t p = 0;
    if (log_evalue(raw, lb, V.totalA, V.nA) <= V.log_cut) { p |= 1; if (V.top_n == 1) atomicMax(&V.bestF[V.gidxB[g.tb]], raw); }
    if (log_evalue(raw, la, V.totalB, V.nB) <= V.log_cut) { p |= 2; if (V.top_n == 1) atomicMax(&V.bestB[V.gidxA[g.ta]], raw); }
    V.pass[x] = p;
}
// top_n > 1: sort keys (query gene, score descending) of one direction's rows; rows that failed the e-value
// cut sort behind every gene
__global__ void k_sel_topn_keys(SelView V, int backward, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    const uint32_t gene = backward ? (uint32_t)V.gidxA[g.ta] : (uint32_t)V.gidxB[g.tb];
    const bool in = V.pass[x] & (backward ? 2 : 1);
    keys[x] = in ? (((uint64_t)gene << 32) | (uint32_t)(0x7fffffff - (g.score2 >> 1))) : EMPTY64;
    vals[x] = (uint32_t)x;
}
// threshold of every gene = score of its top_n-th row (or of its last row when it has fewer)
__global__ void k_sel_topn_thr(const uint64_t* __restrict__ keys, int64_t n, int32_t top_n, int32_t* __restrict__ thr) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = keys[i];
    if (k == EMPTY64) return;
    const uint64_t first = k & 0xFFFFFFFF00000000ull;
    int64_t lo = 0, hi = i;                       // first row of this gene
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < first) lo = mid + 1; else hi = mid;
    }
    const int64_t rank = i - lo;
    const bool last = i + 1 == n || (keys[i + 1] >> 32) != (k >> 32);
    if (rank == top_n - 1 || (last && rank < top_n - 1)) thr[k >> 32] = 0x7fffffff - (int32_t)(uint32_t)k;
}
// membership of HSP x in F (bit0) / Bk (bit1)
__device__ __forceinline__ uint32_t membership(const SelView& V, int64_t x, const Gapped& g, int32_t raw) {
    uint8_t p = V.pass[x];
    uint32_t mem = 0;
    if ((p & 1) && raw >= V.bestF[V.gidxB[g.tb]]) mem |= 1;
    if ((p & 2) && raw >= V.bestB[V.gidxA[g.ta]]) mem |= 2;
    return mem;
}
__device__ __forceinline__ uint64_t pair_key(const SelView& V, const Gapped& g) {
    return ((uint64_t)(uint32_t)V.gidxB[g.tb] << 32) | (uint32_t)V.gidxA[g.ta];
}
// pass 2: register which direction frames contain each (qgene, sgene) pair
__global__ void k_sel_pairs(SelView V) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    uint32_t mem = membership(V, x, g, g.score2 >> 1);
    if (!mem) return;
    uint32_t s = pair_slot(V, pair_key(V, g), true);
    atomicOr(&V.hflags[s], mem);
}
// pass 3: best score per reciprocal pair
__global__ void k_sel_pairbest(SelView V) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    int32_t raw = g.score2 >> 1;
    if (!membership(V, x, g, raw)) return;
    uint32_t s = pair_slot(V, pair_key(V, g), false);
    if (s != 0xFFFFFFFFu && V.hflags[s] == 3) atomicMax(&V.hbest[s], raw);
}
// pass 4: best score per query gene among the pair winners
__global__ void k_sel_genebest(SelView V) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    int32_t raw = g.score2 >> 1;
    if (!membership(V, x, g, raw)) return;
    uint32_t s = pair_slot(V, pair_key(V, g), false);
    if (s != 0xFFFFFFFFu && V.hflags[s] == 3 && V.hbest[s] == raw) atomicMax(&V.bestQ[V.gidxB[g.tb]], raw);
}
// sort key of a table row: qgene asc, sgene asc, forward frame before backward, HSP order
__device__ __forceinline__ uint64_t row_key(const SelView& V, const Gapped& g, int64_t x, uint32_t rev) {
    return ((uint64_t)(uint32_t)V.gidxB[g.tb] << 43) | ((uint64_t)(uint32_t)V.gidxA[g.ta] << 23) |
           ((uint64_t)rev << 22) | (uint64_t)x;
}
// pass 5 (keep="first" only): the first surviving row of each query gene
__global__ void k_sel_first(SelView V) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    int32_t raw = g.score2 >> 1;
    uint32_t mem = membership(V, x, g, raw);
    if (!mem) return;
    uint32_t s = pair_slot(V, pair_key(V, g), false);
    if (s == 0xFFFFFFFFu || V.hflags[s] != 3 || V.hbest[s] != raw || V.bestQ[V.gidxB[g.tb]] != raw) return;
    for (uint32_t rev = 0; rev < 2; rev++)
        if (mem & (1u << rev)) atomicMin(&V.firstkey[V.gidxB[g.tb]], (unsigned long long)row_key(V, g, x, rev));
}
// pass 6: emit surviving rows as (sort key, descriptor = x<<1 | reverse)
__global__ void k_sel_emit(SelView V, int keep_all, uint64_t* __restrict__ keys, uint32_t* __restrict__ desc,
                           unsigned long long* __restrict__ counters /* [0]=rows [1]=nF [2]=nB */) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= V.n) return;
    const Gapped g = V.G[x];
    int32_t raw = g.score2 >> 1;
    uint32_t mem = membership(V, x, g, raw);
    if (!mem) return;
    if (mem & 1) atomicAdd(counters + 1, 1ull);
    if (mem & 2) atomicAdd(counters + 2, 1ull);
    uint32_t s = pair_slot(V, pair_key(V, g), false);
    if (s == 0xFFFFFFFFu || V.hflags[s] != 3 || V.hbest[s] != raw || V.bestQ[V.gidxB[g.tb]] != raw) return;
    for (uint32_t rev = 0; rev < 2; rev++) {
        if (!(mem & (1u << rev))) continue;
        uint64_t k = row_key(V, g, x, rev);
        if (!keep_all && V.firstkey[V.gidxB[g.tb]] != k) continue;
        unsigned long long slot = atomicAdd(counters, 1ull);
        keys[slot] = k;
        desc[slot] = ((uint32_t)x << 1) | rev;
    }
}

// pass 7: materialise the table columns in sorted row order
__global__ void k_sel_fill(const Gapped* __restrict__ G, const uint32_t* __restrict__ desc, int64_t nrows,
                           const int32_t* __restrict__ lenA, const int32_t* __restrict__ lenB,
                           const int32_t* __restrict__ geneA, const int32_t* __restrict__ isoA,
                           const int32_t* __restrict__ geneB, const int32_t* __restrict__ isoB,
                           int32_t* __restrict__ cols) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    uint32_t d = desc[r];
    const Gapped g = G[d >> 1];
    const int rev = d & 1;
    const int32_t la = lenA[g.ta], lb = lenB[g.tb];
    const int32_t p0 = g.minus ? lb - g.b1 : g.b0, p1 = g.minus ? lb - g.b0 : g.b1;  // B plus coords
    int32_t qstart, qend, sstart, send, qlen;
    if (!rev) {  // query = B transcript, subject = A
        qlen = lb; qstart = p0 + 1; qend = p1;
        if (!g.minus) { sstart = g.a0 + 1; send = g.a1; } else { sstart = g.a1; send = g.a0 + 1; }
    } else {     // query = A transcript, subject = B
        qlen = la; qstart = g.a0 + 1; qend = g.a1;
        if (!g.minus) { sstart = p0 + 1; send = p1; } else { sstart = p1; send = p0 + 1; }
    }
#define COL(c) cols[(int64_t)(c) * nrows + r]
    COL(RCQ_COL_LENGTH) = g.nident + g.mism + g.gaps;
    COL(RCQ_COL_MISMATCH) = g.mism; COL(RCQ_COL_GAPOPEN) = g.gapopen;
    COL(RCQ_COL_QSTART) = qstart; COL(RCQ_COL_QEND) = qend; COL(RCQ_COL_SSTART) = sstart; COL(RCQ_COL_SEND) = send;
    COL(RCQ_COL_SCORE) = g.score2 >> 1; COL(RCQ_COL_GAPS) = g.gaps; COL(RCQ_COL_NIDENT) = g.nident;
    // after the reference's rename, qgene/qiso always name the sample-2 transcript
    COL(RCQ_COL_QGENE) = geneB[g.tb]; COL(RCQ_COL_QISO) = isoB[g.tb];
    COL(RCQ_COL_SGENE) = geneA[g.ta]; COL(RCQ_COL_SISO) = isoA[g.ta];
    COL(RCQ_COL_QLEN) = qlen; COL(RCQ_COL_MINUS) = g.minus; COL(RCQ_COL_REVERSE) = rev;
#undef COL
}
