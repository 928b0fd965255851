// search.cu — host orchestration of one unordered pair search (rcq_search_pair) and the
// rcq_table_* entry points.  Replaces, per sample pair, two `blastn` subprocesses plus the
// pandas post-processing of find_homologs.py:82-164,215-302.
#include <cub/cub.cuh>
#include <math.h>
#include <cstdlib>
#include <algorithm>
#include "search.cuh"

#include "seed_kernels.cuh"
#include "gapped_kernels.cuh"
#include "gapped_reg.cuh"
#include "select_kernels.cuh"

struct GappedValid {
    __host__ __device__ bool operator()(const Gapped& g) const { return g.ta >= 0; }
};

static int sort_pairs_u64_u32(rcq_ctx* ctx, uint64_t** keys, uint32_t** vals, int64_t n, int end_bit) {
    if (n <= 1) return 0;
    uint64_t* k2 = nullptr; uint32_t* v2 = nullptr;
    RC(talloc(ctx, &k2, n));
    RC(talloc(ctx, &v2, n));
    cub::DoubleBuffer<uint64_t> dk(*keys, k2);
    cub::DoubleBuffer<uint32_t> dv(*vals, v2);
    size_t bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, dk, dv, (int)n, 0, end_bit, ctx->stream));
    void* ws = nullptr;
    RC(talloc(ctx, (char**)&ws, (int64_t)bytes));
    CU(cub::DeviceRadixSort::SortPairs(ws, bytes, dk, dv, (int)n, 0, end_bit, ctx->stream));
    ctx->launches += (end_bit + 7) / 8 + 1;
    uint64_t* ko = dk.Current(); uint32_t* vo = dv.Current();
    tfree(ctx, ko == *keys ? k2 : *keys);
    tfree(ctx, vo == *vals ? v2 : *vals);
    *keys = ko; *vals = vo;
    return 0;
}

static inline int bits_for(int64_t n) {  // bits needed to hold values 0 .. n-1 (at least 1)
    int b = 1;
    while ((1ll << b) < n) b++;
    return b;
}

// Fallback rows of the gapped kernel: a pool of slots owned by the context, grown when a pair with
// longer transcripts arrives (slots x rows of 6 * (max_len1 + max_len2 + 8) int32, at most ~1 GB).
static int scratch_pool(rcq_ctx* ctx, int64_t per_slot, ScratchPool* sp) {
    if (per_slot > ctx->gs_per_slot) {
        int64_t n_slots = RCQ_SMS * 4;
        while (n_slots > 16 && n_slots * per_slot * 4 > (1ll << 30)) n_slots /= 2;
        CU(cudaStreamSynchronize(ctx->stream));
        if (ctx->gs_rows) CU(cudaFree(ctx->gs_rows));
        if (ctx->gs_locks) CU(cudaFree(ctx->gs_locks));
        ctx->gs_rows = nullptr; ctx->gs_locks = nullptr; ctx->gs_per_slot = 0;
        CU(cudaMalloc((void**)&ctx->gs_rows, (size_t)n_slots * per_slot * 4));
        CU(cudaMalloc((void**)&ctx->gs_locks, (size_t)n_slots * sizeof(unsigned int)));
        CU(cudaMemset(ctx->gs_locks, 0, (size_t)n_slots * sizeof(unsigned int)));
        ctx->gs_per_slot = per_slot;
        ctx->gs_slots = (int32_t)n_slots;
    }
    sp->rows = ctx->gs_rows; sp->per_slot = ctx->gs_per_slot; sp->locks = ctx->gs_locks; sp->n_slots = ctx->gs_slots;
    return 0;
}

static int read_counters(rcq_ctx* ctx, int n) {
    return publish_counters(ctx, ctx->d_ctr, n);
}

// The register-row passes of the gapped stage: LANES x C1 diagonals for every group, LANES x C2 for the groups the
// first pass gave up.  ctr: [0] / [5] read positions, [4] / [6] lengths of the two lists (rcq_search_pair below).
template <int LANES, int C1, int C2, int THREADS, int MINB>
static int launch_gapped_reg(rcq_ctx* ctx, int* occ_slot, const Ungapped* U, const uint64_t* keys, int64_t n_hsp, KeyFmt F,
                             const uint32_t* heads, int64_t n_groups, PairView P, Gapped* G, unsigned long long* ctr,
                             uint32_t* list1, uint32_t* list2) {
    auto k1 = k_gapped_reg<LANES, C1, THREADS, MINB>;
    auto k2 = k_gapped_reg<LANES, C2, THREADS, MINB>;
    const size_t smem = gr_smem_bytes<LANES, THREADS>();
    if (!*occ_slot) {
        CU(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CU(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int nb = 0;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k1, THREADS, smem));
        *occ_slot = nb > 0 ? nb : 1;
        if (const char* e = getenv("RCQ_GAPPED_BPS")) *occ_slot = std::max(1, std::min(*occ_slot, atoi(e)));
    }
    const int blocks = RCQ_SMS * *occ_slot;   // persistent engines, groups handed out by an atomic counter
    LAUNCH(ctx, k1, blocks, THREADS, smem, U, keys, n_hsp, F, heads, n_groups, P, G,
           (const uint32_t*)nullptr, (const unsigned long long*)nullptr, ctr, list1, ctr + 4, ctr);
    LAUNCH(ctx, k2, blocks, THREADS, smem, U, keys, n_hsp, F, heads, n_groups, P, G,
           (const uint32_t*)list1, (const unsigned long long*)(ctr + 4), ctr + 5, list2, ctr + 6, ctr);
    return RCQ_OK;
}

extern "C" int rcq_search_pair(rcq_ctx* ctx, const rcq_sample* s1, const rcq_sample* s2, double evalue,
                               int32_t top_n, int32_t keep_all, rcq_table** out, rcq_search_stats* stats) {
    if (!ctx || !s1 || !s2 || !out) return rcq_fail(RCQ_ERR_ARG, "rcq_search_pair: bad argument");
    if (top_n < 1) return rcq_fail(RCQ_ERR_ARG, "rcq_search_pair: top_n must be >= 1");
    if (s1->n_genes >= (1 << 20) || s2->n_genes >= (1 << 20))
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: more than 2^20 genes in a sample");
    CU(cudaSetDevice(ctx->device));
    RC(arena_reset(ctx));
    rcq_search_stats st;
    memset(&st, 0, sizeof st);
    PairView P;
    P.Af = {s1->fwd, s1->n_masked ? s1->mask : nullptr};
    P.Ar = {s1->rc, s1->n_masked ? s1->mask_rc : nullptr};
    P.Bf = {s2->fwd, s2->n_masked ? s2->mask : nullptr};
    P.Br = {s2->rc, s2->n_masked ? s2->mask_rc : nullptr};
    P.offA = s1->off; P.lenA = s1->len; P.offB = s2->off; P.lenB = s2->len;
    P.nA = s1->n_seq; P.nB = s2->n_seq; P.selA = s1->sel; P.selB = s2->sel;
    P.blkA = s1->blk2tx; P.blkB = s2->blk2tx; P.olA = s1->ol; P.olB = s2->ol;
    P.smA = s1->smask; P.smB = s2->smask; P.smBr = s2->smask_rc;
    unsigned long long* ctr = (unsigned long long*)ctx->d_ctr;
    // the table handle reaches the caller only on success: every error return below frees it
    struct TableGuard {
        rcq_table* t;
        ~TableGuard() { if (t) { dfree(t->ctx, t->cols); delete t; } }
    } guard{new rcq_table()};
    rcq_table* T = guard.t;
    T->ctx = ctx;
    *out = nullptr;

    // ---- seeds: probe (index join) -> candidate list -> verify
    Seed* seeds = nullptr;
    uint2* cand = nullptr;
    int64_t n_seeds = 0, cap = std::max<int64_t>(1 << 20, 48ll * std::max(s1->n_seq, s2->n_seq));
    int64_t ccap = std::max<int64_t>(1 << 20, s1->n_index / 2);
    bool seeded = !(s1->n_index && s2->n_index);
    for (int attempt = 0; attempt < 3 && !seeded; attempt++) {
        RC(talloc(ctx, &seeds, cap));
        RC(talloc(ctx, &cand, ccap));
        CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
        stage_tic(ctx, 1);
        stage_tic(ctx, 7);
        LAUNCH(ctx, k_seed_probe, grid_for(s1->n_index, PROBE_TA), PROBE_THREADS, 0, s1->index, s1->fctx, s1->n_index, s2->index,
               s2->fctx, s2->dir, cand, ccap, ctr);
        stage_toc(ctx, 7);
        {   // candidates into position order of sample 1 (counting sort on the top bits; the count stays on the device)
            uint2* cand2 = nullptr; uint32_t* hist = nullptr;
            RC(talloc(ctx, &cand2, ccap));
            RC(talloc(ctx, &hist, (int64_t)CAND_BUCKETS * CAND_PAD));
            const int shift = std::max(0, bits_for(s1->arena) - CAND_BUCKET_BITS);
            CU(cudaMemsetAsync(hist, 0, (size_t)CAND_BUCKETS * CAND_PAD * sizeof(uint32_t), ctx->stream));
            LAUNCH(ctx, k_cand_hist, RCQ_SMS * 4, 256, 0, cand, ccap, ctr, shift, hist);
            LAUNCH(ctx, k_cand_scan, 1, 1024, 0, hist);
            LAUNCH(ctx, k_cand_scatter, RCQ_SMS * 8, 256, 0, cand, ccap, ctr, shift, hist, cand2);
            tfree(ctx, cand); tfree(ctx, hist);
            cand = cand2;
        }
        LAUNCH(ctx, k_seed_verify, RCQ_SMS * 16, 256, 0, cand, ccap, P, seeds, cap, ctr);
        stage_toc(ctx, 1);
        RC(read_counters(ctx, 3));
        stage_collect(ctx);
        n_seeds = ctx->h_pin[0];
        st.n_verified = ctx->h_pin[1];
        st.n_candidates = ctx->h_pin[2];
        tfree(ctx, cand); cand = nullptr;
        if (st.n_verified >= (1ll << 31) || n_seeds >= (1ll << 31))
            return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: more than 2^31 seed candidates in one pair - low-complexity "
                            "stretches (poly-A tails, microsatellites) shared by many transcripts seed against each other; "
                            "the DUST filter that keeps them out of seeding is off for these samples (rcq_ctx_set_dust) "
                            "or does not cover this repeat");
        if (n_seeds <= cap && st.n_verified <= ccap) { seeded = true; break; }
        tfree(ctx, seeds); seeds = nullptr;
        cap = std::max(cap, n_seeds);
        ccap = std::max(ccap, st.n_verified);
        n_seeds = 0;
    }
    if (!seeded) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: the seed lists kept growing between attempts");
    st.n_seeds = n_seeds;
    if (n_seeds >= (1ll << 31)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: too many seeds");
    int64_t n_gapped = 0;
    Gapped* Gd = nullptr;
    if (n_seeds > 0) {
        // ---- ungapped HSPs under the diagonal rule: first HSP of every diagonal through the diagonal
        // table, the seeds that start beyond it through the sort path (seed_kernels.cuh)
        if (s1->n_seq >= (1 << 21) || s2->n_seq >= (1 << 21))
            return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: more than 2^21 transcripts in a sample");
        KeyFmt F;
        F.bias = s1->max_len;
        F.d_bits = bits_for((int64_t)s1->max_len + s2->max_len + 1);
        F.tb_bits = bits_for(s2->n_seq);
        F.total_bits = F.d_bits + 1 + F.tb_bits + bits_for(s1->n_seq);
        if (F.total_bits > 63) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: seed key does not fit 63 bits");
        DiagTable T;
        int64_t tcap = 1024;
        while (tcap < 2 * n_seeds) tcap <<= 1;
        T.mask = (uint32_t)(tcap - 1);
        RC(talloc(ctx, &T.keys, 2 * tcap));
        T.vals = T.keys + tcap;
        uint32_t* slots = nullptr;
        Ungapped* U1 = nullptr;
        Seed* surv = nullptr;
        RC(talloc(ctx, &slots, n_seeds));
        RC(talloc(ctx, &U1, n_seeds));
        RC(talloc(ctx, &surv, n_seeds));
        CU(cudaMemsetAsync(T.keys, 0xFF, 2 * tcap * sizeof(unsigned long long), ctx->stream));
        CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
        stage_tic(ctx, 2);
        LAUNCH(ctx, k_diag_insert, grid_for(n_seeds, 256), 256, 0, seeds, n_seeds, F, T, slots, ctr);
        LAUNCH(ctx, k_diag_first, grid_for(2 * n_seeds, 128), 128, 0, seeds, slots, T, P, U1, ctr);
        LAUNCH(ctx, k_diag_survivors, grid_for(n_seeds, 256), 256, 0, seeds, n_seeds, F, T, surv, ctr);
        stage_toc(ctx, 2);
        RC(read_counters(ctx, 10));
        stage_collect(ctx);
        const int64_t n_diag = ctx->h_pin[6], n_surv = ctx->h_pin[9];
        st.ungapped_cells = ctx->h_pin[4];
        tfree(ctx, T.keys); tfree(ctx, slots);
        tfree(ctx, seeds); seeds = nullptr;
        const int64_t n_hsp = n_diag + n_surv;
        Ungapped* Uall = nullptr;
        RC(talloc(ctx, &Uall, n_hsp));
        CU(cudaMemcpyAsync(Uall, U1, n_diag * sizeof(Ungapped), cudaMemcpyDeviceToDevice, ctx->stream));
        tfree(ctx, U1);
        CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
        if (n_surv > 0) {
            // survivors: sorted by (tx1, tx2, strand, diagonal), thread per diagonal continues the rule
            uint64_t* keys = nullptr; uint32_t* order = nullptr; uint32_t* heads = nullptr; uint32_t* dheads = nullptr;
            RC(talloc(ctx, &keys, n_surv));
            RC(talloc(ctx, &order, n_surv));
            RC(talloc(ctx, &heads, n_surv));
            RC(talloc(ctx, &dheads, n_surv));
            CU(cudaMemsetAsync(Uall + n_diag, 0xFF, n_surv * sizeof(Ungapped), ctx->stream));
            stage_tic(ctx, 3);
            LAUNCH(ctx, k_seed_keys, grid_for(n_surv, 256), 256, 0, surv, n_surv, F, keys, order);
            RC(sort_pairs_u64_u32(ctx, &keys, &order, n_surv, F.total_bits));
            LAUNCH(ctx, k_seed_heads, grid_for(n_surv, 256), 256, 0, keys, n_surv, F, dheads, heads, ctr);
            stage_toc(ctx, 3);
            RC(read_counters(ctx, 7));
            stage_collect(ctx);
            const int64_t n_diag2 = ctx->h_pin[6];
            LAUNCH(ctx, k_ungapped_diag, grid_for(n_diag2, 128), 128, 0, surv, order, keys, n_surv, dheads, n_diag2, P,
                   Uall + n_diag, ctr);
            RC(read_counters(ctx, 5));
            st.ungapped_cells += ctx->h_pin[4];
            CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
            tfree(ctx, keys); tfree(ctx, order); tfree(ctx, heads); tfree(ctx, dheads);
        }
        tfree(ctx, surv);
        // ---- HSPs grouped by (tx1, tx2, strand)
        uint64_t* keys = nullptr; uint32_t* order = nullptr; uint32_t* heads = nullptr;
        Ungapped* U = nullptr;
        RC(talloc(ctx, &keys, n_hsp));
        RC(talloc(ctx, &order, n_hsp));
        RC(talloc(ctx, &heads, n_hsp));
        RC(talloc(ctx, &U, n_hsp));
        stage_tic(ctx, 3);
        LAUNCH(ctx, k_hsp_keys, grid_for(n_hsp, 256), 256, 0, Uall, n_hsp, F, keys, order, ctr);
        RC(sort_pairs_u64_u32(ctx, &keys, &order, n_hsp, F.total_bits - F.d_bits + 1));
        LAUNCH(ctx, k_hsp_gather, grid_for(n_hsp, 256), 256, 0, Uall, order, keys, n_hsp, F, U, heads, ctr);
        stage_toc(ctx, 3);
        RC(read_counters(ctx, 8));
        stage_collect(ctx);
        const int64_t n_groups = ctx->h_pin[5];
        st.n_ungapped = ctx->h_pin[7];
        tfree(ctx, Uall); tfree(ctx, order);
        const int64_t n_slots = n_hsp;   // gapped HSP slots: group g writes into the range of its ungapped HSPs
        // ---- gapped extension, warp per group
        Gapped* G = nullptr;
        RC(talloc(ctx, &G, n_slots));
        CU(cudaMemsetAsync(G, 0xFF, n_slots * sizeof(Gapped), ctx->stream));
        const int threads = 128;
        const bool masked = s1->n_masked || s2->n_masked;
        int* occ = ctx->gapped_occ;   // resident blocks per SM: [0] k_gapped_reg<8>, [1] k_gapped<true>, [2] k_gapped_reg<16> (per context: contexts run on their own threads)
        if (masked && !occ[1]) {
            int nb = 0;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_gapped<true>, threads, 0));
            occ[1] = nb > 0 ? nb : 1;
        }
        ScratchPool sp;
        RC(scratch_pool(ctx, 6ll * ((int64_t)s1->max_len + s2->max_len + 8), &sp));
        CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
        stage_tic(ctx, 4);
        if (masked) {
            // samples with N bases: the general path (mask words consulted in every comparison)
            LAUNCH(ctx, k_gapped<true>, RCQ_SMS * occ[1], threads, 0, U, keys, n_hsp, F, heads, n_groups, P, G, sp, ctr,
                   (const uint32_t*)nullptr, (const unsigned long long*)nullptr, (unsigned long long*)nullptr);
        } else {
            // engines of 8 lanes on register rows: 48-diagonal windows for every group, 64-diagonal windows for the
            // groups whose band outgrew 48 (a few per cent), and the wide kernel on global rows for what is left
            // (bands beyond 64, transcripts beyond the staged length).  ctr: [4] / [6] lengths of the two lists,
            // [5] / [7] their read positions.  RCQ_GAPPED_VARIANT=16 selects half-warp engines (measurement only).
            uint32_t *list1 = nullptr, *list2 = nullptr;
            RC(talloc(ctx, &list1, n_groups));
            RC(talloc(ctx, &list2, n_groups));
            const char* variant = getenv("RCQ_GAPPED_VARIANT");
            if (variant && atoi(variant) == 16)
                RC((launch_gapped_reg<16, 3, 4, 128, 4>(ctx, &occ[2], U, keys, n_hsp, F, heads, n_groups, P, G, ctr, list1, list2)));
            else
                RC((launch_gapped_reg<8, 6, 8, 128, 4>(ctx, &occ[0], U, keys, n_hsp, F, heads, n_groups, P, G, ctr, list1, list2)));
            LAUNCH(ctx, k_gapped<false>, RCQ_SMS * 2, threads, 0, U, keys, n_hsp, F, heads, n_groups, P, G, sp, ctr,
                   (const uint32_t*)list2, (const unsigned long long*)(ctr + 6), ctr + 7);
            tfree(ctx, list1); tfree(ctx, list2);
        }
        stage_toc(ctx, 4);
        tfree(ctx, U); tfree(ctx, keys); tfree(ctx, heads);
        // ---- compact (stable: group order, then processing order inside the group)
        RC(talloc(ctx, &Gd, n_slots));
        size_t bytes = 0;
        CU(cub::DeviceSelect::If(nullptr, bytes, G, Gd, (int*)(ctr + 8), (int)n_slots, GappedValid(), ctx->stream));
        void* ws = nullptr;
        RC(talloc(ctx, (char**)&ws, (int64_t)bytes));
        CU(cub::DeviceSelect::If(ws, bytes, G, Gd, (int*)(ctr + 8), (int)n_slots, GappedValid(), ctx->stream));
        ctx->launches += 2;
            tfree(ctx, G);
        RC(read_counters(ctx, 9));
        st.dp_cells = ctx->h_pin[1];
        st.band_retries = ctx->h_pin[2];
        n_gapped = (int64_t)(int32_t)(ctx->h_pin[8] & 0xFFFFFFFF);
        st.n_gapped = n_gapped;
    } else {
        tfree(ctx, seeds);
    }
    // debug copy of the HSP list (test hook, off unless rcq_debug_keep_hsps(ctx, 1): cudaMalloc/cudaFree
    // synchronise the whole device and would serialise the streams of the engine)
    if (ctx->dbg_keep) {
        if (ctx->dbg_hsp) { cudaFree(ctx->dbg_hsp); ctx->dbg_hsp = nullptr; }
        ctx->dbg_n = n_gapped;
        if (n_gapped) {
            CU(cudaMalloc((void**)&ctx->dbg_hsp, n_gapped * sizeof(Gapped)));
            CU(cudaMemcpyAsync(ctx->dbg_hsp, Gd, n_gapped * sizeof(Gapped), cudaMemcpyDeviceToDevice, ctx->stream));
        }
    }
    if (n_gapped >= (1 << 22)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_search_pair: more than 2^22 gapped HSPs in one pair");
    // ---- selection + table
    int64_t n_rows = 0;
    if (n_gapped > 0) {
        SelView V;
        V.G = Gd; V.n = n_gapped;
        V.lenA = s1->len; V.lenB = s2->len; V.gidxA = s1->gidx; V.gidxB = s2->gidx;
        V.totalA = s1->total; V.totalB = s2->total; V.nA = s1->n_seq; V.nB = s2->n_seq;
        V.log_cut = evalue > 0 ? log(evalue) : -INFINITY;
        V.top_n = top_n;
        int64_t hcap = 1024;
        while (hcap < 4 * n_gapped) hcap <<= 1;
        V.hmask = (uint32_t)(hcap - 1);
        RC(talloc(ctx, &V.pass, n_gapped));
        RC(talloc(ctx, &V.bestF, s2->n_genes));
        RC(talloc(ctx, &V.bestB, s1->n_genes));
        RC(talloc(ctx, &V.bestQ, s2->n_genes));
        RC(talloc(ctx, &V.firstkey, s2->n_genes));
        RC(talloc(ctx, &V.hkeys, hcap));
        RC(talloc(ctx, &V.hflags, hcap));
        RC(talloc(ctx, &V.hbest, hcap));
        CU(cudaMemsetAsync(V.bestF, 0xFF, s2->n_genes * 4, ctx->stream));
        CU(cudaMemsetAsync(V.bestB, 0xFF, s1->n_genes * 4, ctx->stream));
        CU(cudaMemsetAsync(V.bestQ, 0xFF, s2->n_genes * 4, ctx->stream));
        CU(cudaMemsetAsync(V.firstkey, 0xFF, s2->n_genes * 8, ctx->stream));
        CU(cudaMemsetAsync(V.hkeys, 0xFF, hcap * 8, ctx->stream));
        CU(cudaMemsetAsync(V.hflags, 0, hcap * 4, ctx->stream));
        CU(cudaMemsetAsync(V.hbest, 0xFF, hcap * 4, ctx->stream));
        CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
        uint64_t* rkeys = nullptr; uint32_t* rdesc = nullptr;
        RC(talloc(ctx, &rkeys, 2 * n_gapped));
        RC(talloc(ctx, &rdesc, 2 * n_gapped));
        int g = grid_for(n_gapped, 256);
        stage_tic(ctx, 5);
        LAUNCH(ctx, k_sel_best, g, 256, 0, V);
        if (top_n > 1) {
            // n-largest per query gene (find_homologs.py:135-164 with n > 1): rows sorted by (gene, score desc),
            // the n-th row of every gene is its threshold.  rkeys/rdesc are free until k_sel_emit.
            for (int backward = 0; backward < 2; backward++) {
                uint64_t* tk = rkeys; uint32_t* tv = rdesc;
                LAUNCH(ctx, k_sel_topn_keys, g, 256, 0, V, backward, tk, tv);
                RC(sort_pairs_u64_u32(ctx, &tk, &tv, n_gapped, 64));
                LAUNCH(ctx, k_sel_topn_thr, g, 256, 0, tk, n_gapped, top_n, backward ? V.bestB : V.bestF);
            }
        }
        LAUNCH(ctx, k_sel_pairs, g, 256, 0, V);
        LAUNCH(ctx, k_sel_pairbest, g, 256, 0, V);
        LAUNCH(ctx, k_sel_genebest, g, 256, 0, V);
        if (!keep_all) LAUNCH(ctx, k_sel_first, g, 256, 0, V);
        LAUNCH(ctx, k_sel_emit, g, 256, 0, V, (int)keep_all, rkeys, rdesc, ctr);
        RC(read_counters(ctx, 3));
        n_rows = ctx->h_pin[0];
        st.n_fwd = ctx->h_pin[1];
        st.n_bwd = ctx->h_pin[2];
        if (n_rows > 0) {
            RC(sort_pairs_u64_u32(ctx, &rkeys, &rdesc, n_rows, 64));
            RC(dalloc(ctx, &T->cols, (int64_t)RCQ_NCOLS * n_rows));
            LAUNCH(ctx, k_sel_fill, grid_for(n_rows, 256), 256, 0, Gd, rdesc, n_rows, s1->len, s2->len,
                   s1->gene, s1->iso, s2->gene, s2->iso, T->cols);
        }
        stage_toc(ctx, 5);
        tfree(ctx, rkeys); tfree(ctx, rdesc);
        tfree(ctx, V.pass); tfree(ctx, V.bestF); tfree(ctx, V.bestB); tfree(ctx, V.bestQ);
        tfree(ctx, V.firstkey); tfree(ctx, V.hkeys); tfree(ctx, V.hflags); tfree(ctx, V.hbest);
    }
    tfree(ctx, Gd);
    T->n = n_rows;
    st.n_rows = n_rows;
    CU(cudaStreamSynchronize(ctx->stream));
    stage_collect(ctx);
    if (stats) *stats = st;
    guard.t = nullptr;
    *out = T;
    return 0;
}

extern "C" int64_t rcq_table_rows(const rcq_table* t) { return t ? t->n : 0; }

extern "C" void rcq_table_free(rcq_table* t) {
    if (!t) return;
    cudaSetDevice(t->ctx->device);
    dfree(t->ctx, t->cols);
    delete t;
}

extern "C" int rcq_table_download(const rcq_table* t, rcq_table_cols* h) {
    if (!t || !h) return rcq_fail(RCQ_ERR_ARG, "rcq_table_download: bad argument");
    if (t->n == 0) return 0;
    rcq_ctx* ctx = t->ctx;
    CU(cudaSetDevice(ctx->device));
    // one DMA transfer into the context's page-locked staging (a pageable destination costs a staged copy and a page
    // fault per 4 KB: 13 ms per 100,000-row table with six streams downloading at once), then plain copies
    const size_t words = (size_t)RCQ_NCOLS * t->n;
    RC(stage_reserve(ctx, words * 4));
    CU(cudaMemcpyAsync(ctx->h_stage, t->cols, words * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const int32_t* buf = (const int32_t*)ctx->h_stage;
    int32_t* dst[15] = {h->length, h->mismatch, h->gapopen, h->qstart, h->qend, h->sstart, h->send, h->score,
                        h->gaps, h->nident, h->qgene, h->qiso, h->sgene, h->siso, h->qlen};
    for (int c = 0; c < 15; c++)
        if (dst[c]) memcpy(dst[c], buf + (size_t)c * t->n, t->n * 4);
    for (int64_t r = 0; r < t->n; r++) {
        if (h->minus) h->minus[r] = (uint8_t)buf[(size_t)RCQ_COL_MINUS * t->n + r];
        if (h->reverse) h->reverse[r] = (uint8_t)buf[(size_t)RCQ_COL_REVERSE * t->n + r];
    }
    return 0;
}

extern "C" int rcq_debug_keep_hsps(rcq_ctx* ctx, int on) {
    if (!ctx) return 0;
    const int before = ctx->dbg_keep ? 1 : 0;
    ctx->dbg_keep = on;
    return before;
}

extern "C" int64_t rcq_debug_last_hsps(rcq_ctx* ctx, int32_t* out, int64_t cap) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    int64_t n = ctx->dbg_n < cap ? ctx->dbg_n : cap;
    if (out && n > 0) cudaMemcpy(out, ctx->dbg_hsp, n * sizeof(Gapped), cudaMemcpyDeviceToHost);
    return ctx->dbg_n;
}
