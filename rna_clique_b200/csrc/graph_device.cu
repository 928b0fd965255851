// graph_device.cu — graph.pkl from the DEVICE-RESIDENT gene-matches tables: everything build_graph
// (build_graph.py:40-68) decides — node numbering by first appearance, one edge per distinct (sgene, qgene)
// row, adjacency lists in insertion order, which visit of an edge creates its attribute dict — is computed on the
// GPU from the tables' gene columns; the host only turns the resulting arrays into pickle bytes (rcq_graph_emit,
// host_graph.cu).  Same bytes as rcq_graph_pickle_groups on the downloaded columns (tests/test_graph_pickle.py on
// the CPU, tests/test_api_gpu.py on the device); at BASELINE configs[2] (435 tables, 43.5 million rows) the host
// preparation took 0.6 s on 16 cores, the kernels below take milliseconds.
//
// Requirements (the caller falls back to the host writer when one fails, RCQ_ERR_ARG / RCQ_ERR_CAPACITY):
// tables as rcq_search_pair writes them (rows ordered by (qgene, sgene)), no two tables joining the same two
// samples, fewer than 2^31 rows, sum over samples of (largest gene id + 1) below 2^28.
// HBM-bound streaming / sorting kernels; algorithmic bytes ~ 60 B per row.
#include <cub/cub.cuh>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "search.cuh"
#include "host_graph.h"

namespace {
struct GTables {                       // device arrays describing the tables
    const int64_t* row_off;            // [n_tables + 1]
    const int32_t* const* cols;        // [n_tables] column blocks (17 column-major int32 arrays)
    const int32_t* ts;                 // [n_tables] sample index of the sgene column
    const int32_t* tq;                 // [n_tables] sample index of the qgene column
    int32_t n_tables;
    int64_t n_rows;
};
__device__ __forceinline__ int32_t g_table_of(const int64_t* __restrict__ off, int32_t n, int64_t r) {
    int32_t lo = 0, hi = n - 1;
    while (lo < hi) {
        const int32_t mid = (lo + hi + 1) >> 1;
        if (__ldg(off + mid) <= r) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ void g_row(const GTables& T, int64_t r, int32_t* t, int32_t* sg, int32_t* qg, bool* dup, bool* unsorted) {
    const int32_t tt = g_table_of(T.row_off, T.n_tables, r);
    const int64_t lo = __ldg(T.row_off + tt), tn = __ldg(T.row_off + tt + 1) - lo, x = r - lo;
    const int32_t* c = T.cols[tt];
    const int32_t s = __ldg(c + RCQ_COL_SGENE * tn + x), q = __ldg(c + RCQ_COL_QGENE * tn + x);
    *t = tt; *sg = s; *qg = q;
    if (dup) {
        *dup = false; *unsorted = false;
        if (x > 0) {
            const int32_t ps = __ldg(c + RCQ_COL_SGENE * tn + x - 1), pq = __ldg(c + RCQ_COL_QGENE * tn + x - 1);
            *dup = ps == s && pq == q;
            *unsorted = q < pq || (q == pq && s < ps);
        }
    }
}
// largest gene id of every sample (and a check for negative ids)
__global__ void k_g_maxgene(GTables T, int32_t* __restrict__ max_gene, unsigned long long* __restrict__ flags) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= T.n_rows) return;
    int32_t t, s, q;
    g_row(T, r, &t, &s, &q, nullptr, nullptr);
    if (s < 0 || q < 0) { atomicOr(flags, 1ull); return; }
    atomicMax(max_gene + __ldg(T.ts + t), s);
    atomicMax(max_gene + __ldg(T.tq + t), q);
}
// first appearance of every (sample, gene): the smallest "row << 1 | side" that names it
__global__ void k_g_first(GTables T, const int64_t* __restrict__ base, uint32_t* __restrict__ first) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= T.n_rows) return;
    int32_t t, s, q;
    g_row(T, r, &t, &s, &q, nullptr, nullptr);
    atomicMin(first + __ldg(base + __ldg(T.ts + t)) + s, (uint32_t)(2 * r));
    atomicMin(first + __ldg(base + __ldg(T.tq + t)) + q, (uint32_t)(2 * r + 1));
}
struct NotInf { __host__ __device__ bool operator()(uint32_t x) const { return x != 0xFFFFFFFFu; } };
// nodes in first-appearance order: slot -> node id, node id -> (sample, gene)
__global__ void k_g_nodes(const uint32_t* __restrict__ slot_sorted, int64_t n_nodes, const int64_t* __restrict__ base, int32_t n_samples,
                          int32_t* __restrict__ slot_id, int32_t* __restrict__ node_sample, int32_t* __restrict__ node_gene) {
    const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n_nodes) return;
    const int64_t a = slot_sorted[x];
    int32_t lo = 0, hi = n_samples - 1;
    while (lo < hi) {
        const int32_t mid = (lo + hi + 1) >> 1;
        if (__ldg(base + mid) <= a) lo = mid; else hi = mid - 1;
    }
    slot_id[a] = (int32_t)x;
    node_sample[x] = lo; node_gene[x] = (int32_t)(a - __ldg(base + lo));
}
// end points of every row as node ids (u << 32 | v) and whether the row is the first of its edge
__global__ void k_g_rows(GTables T, const int64_t* __restrict__ base, const int32_t* __restrict__ slot_id,
                         uint64_t* __restrict__ uv, uint8_t* __restrict__ keep, unsigned long long* __restrict__ flags) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= T.n_rows) return;
    int32_t t, s, q;
    bool dup, unsorted;
    g_row(T, r, &t, &s, &q, &dup, &unsorted);
    if (unsorted) atomicOr(flags, 2ull);
    const uint32_t u = (uint32_t)slot_id[__ldg(base + __ldg(T.ts + t)) + s], v = (uint32_t)slot_id[__ldg(base + __ldg(T.tq + t)) + q];
    uv[r] = ((uint64_t)u << 32) | v;
    keep[r] = dup ? 0 : 1;
}
// two adjacency entries per edge: key = the node that owns the entry, value = edge << 1 | side (a self-loop is
// listed once: its second entry gets a key behind every node)
__global__ void k_g_entries(const uint64_t* __restrict__ edges, int64_t n_edges, uint32_t n_nodes, uint32_t* __restrict__ key,
                            uint32_t* __restrict__ val, unsigned long long* __restrict__ flags /* [1] += self-loops */) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const uint32_t u = (uint32_t)(edges[e] >> 32), v = (uint32_t)edges[e];
    key[2 * e] = u; val[2 * e] = (uint32_t)(2 * e);
    key[2 * e + 1] = (v == u) ? n_nodes : v; val[2 * e + 1] = (uint32_t)(2 * e + 1);
    if (v == u) atomicAdd(flags + 1, 1ull);
}
// the sorted entries as CSR: neighbour of every entry, start of every node's list
__global__ void k_g_csr(const uint32_t* __restrict__ key, const uint32_t* __restrict__ val, const uint64_t* __restrict__ edges,
                        int64_t m, int64_t n_nodes, int32_t* __restrict__ nbr, int64_t* __restrict__ start) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= m) return;
    const uint32_t k = key[a], x = val[a];
    const uint64_t e = edges[x >> 1];
    nbr[a] = (int32_t)((x & 1u) ? (uint32_t)(e >> 32) : (uint32_t)e);
    if (a == 0 || key[a - 1] != k) start[k] = a;
    if (a == m - 1) start[n_nodes] = m;
}
// first visits in every node's list (entries whose neighbour is not an earlier node)
__global__ void k_g_first_visits(const int64_t* __restrict__ start, const int32_t* __restrict__ nbr, int64_t n_nodes, uint32_t* __restrict__ cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    uint32_t c = 0;
    for (int64_t a = start[i]; a < start[i + 1]; a++) c += nbr[a] >= i;
    cnt[i] = c;
}
// memo number (relative) of every edge's attribute dict, handed out at the first visit ...
__global__ void k_g_edge_memo(const int64_t* __restrict__ start, const int32_t* __restrict__ nbr, const uint32_t* __restrict__ val,
                              const uint32_t* __restrict__ before, int64_t n_nodes, uint32_t* __restrict__ edge_rel) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    uint32_t m = before[i];
    for (int64_t a = start[i]; a < start[i + 1]; a++) if (nbr[a] >= i) edge_rel[val[a] >> 1] = m++;
}
// ... and looked up at the second
__global__ void k_g_entry_memo(const uint32_t* __restrict__ key, const int32_t* __restrict__ nbr, const uint32_t* __restrict__ val,
                               const uint32_t* __restrict__ edge_rel, int64_t m, uint32_t* __restrict__ erel) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= m) return;
    erel[a] = ((uint32_t)nbr[a] >= key[a]) ? RCQ_GRAPH_FIRST_VISIT : edge_rel[val[a] >> 1];
}
}  // namespace

namespace {
struct DevBufs {                 // pool allocations of one call, released together
    rcq_ctx* ctx;
    std::vector<void*> v;
    explicit DevBufs(rcq_ctx* c) : ctx(c) {}
    template <typename T> int get(T** p, int64_t n) { RC(dalloc(ctx, p, n)); v.push_back((void*)*p); return 0; }
    ~DevBufs() { for (void* p : v) cudaFreeAsync(p, ctx->stream); }
};
template <typename T>
static int to_host(rcq_ctx* ctx, std::vector<T>& h, const T* d, int64_t n) {
    h.resize((size_t)n);
    if (n) CU(cudaMemcpyAsync(h.data(), d, (size_t)n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}
static inline int g_bits_for(int64_t n) { int b = 1; while ((1ll << b) < n) b++; return b; }
}  // namespace

extern "C" int rcq_graph_pickle_tables(rcq_ctx* ctx, const char* path, int32_t n_samples, const char* const* sample_names,
                                       int32_t n_tables, const rcq_table* const* tables, const int32_t* t_s, const int32_t* t_q,
                                       int64_t counts[3]) {
    if (!ctx || !path || n_samples < 0 || n_tables < 0 || (n_samples && !sample_names) || (n_tables && (!tables || !t_s || !t_q)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_tables: bad argument");
    try {
        CU(cudaSetDevice(ctx->device));
        // ---- what the host decides: row offsets, no two tables for one pair of samples
        std::vector<int64_t> meta((size_t)2 * n_tables + 1, 0);
        std::vector<uint64_t> pk((size_t)n_tables);
        for (int32_t t = 0; t < n_tables; t++) {
            if (!tables[t] || t_s[t] < 0 || t_s[t] >= n_samples || t_q[t] < 0 || t_q[t] >= n_samples)
                return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_tables: bad table or sample index");
            meta[t + 1] = meta[t] + tables[t]->n;
            meta[n_tables + 1 + t] = (int64_t)(uintptr_t)tables[t]->cols;
            const uint32_t a = (uint32_t)std::min(t_s[t], t_q[t]), b = (uint32_t)std::max(t_s[t], t_q[t]);
            pk[t] = ((uint64_t)a << 32) | b;
        }
        std::sort(pk.begin(), pk.end());
        if (std::adjacent_find(pk.begin(), pk.end()) != pk.end())
            return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_tables: two tables join the same two samples (use rcq_graph_pickle_groups)");
        const int64_t R = meta[n_tables];
        if (R >= (1ll << 31)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle_tables: 2^31 rows or more");
        static const bool trace = getenv("RCQ_TRACE") != nullptr;
        auto T0 = std::chrono::steady_clock::now();
        auto lap = [&](const char* what) {
            if (!trace) return;
            auto T1 = std::chrono::steady_clock::now();
            fprintf(stderr, "[rcq_graph_pickle] %-14s %.3f s\n", what, std::chrono::duration<double>(T1 - T0).count());
            T0 = T1;
        };
        const int64_t zero_start[1] = {0};
        const int32_t *p_ns = nullptr, *p_ng = nullptr, *p_nb = nullptr;
        const int64_t* p_st = zero_start;
        const uint32_t* p_er = nullptr;
        int64_t n_nodes = 0, n_edges = 0;
        if (R > 0) {
            DevBufs D(ctx);
            cudaStream_t st = ctx->stream;
            int64_t* d_meta = nullptr; int32_t* d_tsq = nullptr;
            RC(D.get(&d_meta, (int64_t)meta.size()));
            RC(D.get(&d_tsq, 2ll * n_tables));
            CU(cudaMemcpyAsync(d_meta, meta.data(), meta.size() * 8, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(d_tsq, t_s, (size_t)n_tables * 4, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(d_tsq + n_tables, t_q, (size_t)n_tables * 4, cudaMemcpyHostToDevice, st));
            GTables T;
            T.row_off = d_meta; T.cols = (const int32_t* const*)(d_meta + n_tables + 1); T.ts = d_tsq; T.tq = d_tsq + n_tables;
            T.n_tables = n_tables; T.n_rows = R;
            const int gr = grid_for(R, 256);
            // ---- gene id range of every sample
            int32_t* d_max = nullptr; unsigned long long* d_flags = nullptr;
            RC(D.get(&d_max, (int64_t)n_samples));
            RC(D.get(&d_flags, 4));
            CU(cudaMemsetAsync(d_max, 0xFF, (size_t)n_samples * 4, st));
            CU(cudaMemsetAsync(d_flags, 0, 32, st));
            LAUNCH(ctx, k_g_maxgene, gr, 256, 0, T, d_max, d_flags);
            std::vector<int32_t> max_gene; std::vector<unsigned long long> flags;
            RC(to_host(ctx, max_gene, d_max, n_samples));
            RC(to_host(ctx, flags, (const unsigned long long*)d_flags, 4));
            CU(cudaStreamSynchronize(st));
            if (flags[0] & 1) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_tables: negative gene id");
            std::vector<int64_t> base((size_t)n_samples + 1, 0);
            for (int32_t s = 0; s < n_samples; s++) base[s + 1] = base[s] + max_gene[s] + 1;
            const int64_t span = base[n_samples];
            if (span >= (1ll << 28)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle_tables: gene ids too sparse for dense node tables");
            int64_t* d_base = nullptr;
            RC(D.get(&d_base, (int64_t)base.size()));
            CU(cudaMemcpyAsync(d_base, base.data(), base.size() * 8, cudaMemcpyHostToDevice, st));
            // ---- nodes in first-appearance order
            uint32_t *d_first = nullptr, *d_key = nullptr, *d_slot = nullptr, *d_key2 = nullptr, *d_slot2 = nullptr;
            int64_t* d_num = nullptr;
            RC(D.get(&d_first, span)); RC(D.get(&d_key, span)); RC(D.get(&d_slot, span)); RC(D.get(&d_key2, span)); RC(D.get(&d_slot2, span));
            RC(D.get(&d_num, 2));
            CU(cudaMemsetAsync(d_first, 0xFF, (size_t)span * 4, st));
            LAUNCH(ctx, k_g_first, gr, 256, 0, T, (const int64_t*)d_base, d_first);
            cub::CountingInputIterator<uint32_t> iota(0);
            cub::TransformInputIterator<bool, NotInf, const uint32_t*> present((const uint32_t*)d_first, NotInf());
            size_t b1 = 0, b2 = 0, b3 = 0;
            CU(cub::DeviceSelect::If(nullptr, b1, (const uint32_t*)d_first, d_key, d_num, (int)span, NotInf(), st));
            CU(cub::DeviceSelect::Flagged(nullptr, b2, iota, present, d_slot, d_num + 1, (int)span, st));
            cub::DoubleBuffer<uint32_t> nk(d_key, d_key2), nv(d_slot, d_slot2);
            CU(cub::DeviceRadixSort::SortPairs(nullptr, b3, nk, nv, (int)span, 0, 32, st));
            char* ws = nullptr;
            RC(D.get(&ws, (int64_t)std::max(b1, std::max(b2, b3))));
            CU(cub::DeviceSelect::If(ws, b1, (const uint32_t*)d_first, d_key, d_num, (int)span, NotInf(), st));
            CU(cub::DeviceSelect::Flagged(ws, b2, iota, present, d_slot, d_num + 1, (int)span, st));
            std::vector<int64_t> num;
            RC(to_host(ctx, num, (const int64_t*)d_num, 2));
            CU(cudaStreamSynchronize(st));
            n_nodes = (int64_t)(int32_t)(num[0] & 0xFFFFFFFF);
            if (n_nodes > 1) CU(cub::DeviceRadixSort::SortPairs(ws, b3, nk, nv, (int)n_nodes, 0, 32, st));
            ctx->launches += 8;
            int32_t *d_slot_id = nullptr, *d_ns = nullptr, *d_ng = nullptr;
            RC(D.get(&d_slot_id, span)); RC(D.get(&d_ns, n_nodes)); RC(D.get(&d_ng, n_nodes));
            LAUNCH(ctx, k_g_nodes, grid_for(n_nodes, 256), 256, 0, (const uint32_t*)nv.Current(), n_nodes, (const int64_t*)d_base, n_samples,
                   d_slot_id, d_ns, d_ng);
            // ---- distinct edges in insertion order
            uint64_t *d_uv = nullptr, *d_edges = nullptr; uint8_t* d_keep = nullptr;
            RC(D.get(&d_uv, R)); RC(D.get(&d_edges, R)); RC(D.get(&d_keep, R));
            LAUNCH(ctx, k_g_rows, gr, 256, 0, T, (const int64_t*)d_base, (const int32_t*)d_slot_id, d_uv, d_keep, d_flags);
            size_t b4 = 0;
            CU(cub::DeviceSelect::Flagged(nullptr, b4, (const uint64_t*)d_uv, (const uint8_t*)d_keep, d_edges, d_num, (int)R, st));
            char* ws2 = nullptr;
            RC(D.get(&ws2, (int64_t)b4));
            CU(cub::DeviceSelect::Flagged(ws2, b4, (const uint64_t*)d_uv, (const uint8_t*)d_keep, d_edges, d_num, (int)R, st));
            RC(to_host(ctx, num, (const int64_t*)d_num, 1));
            RC(to_host(ctx, flags, (const unsigned long long*)d_flags, 4));
            CU(cudaStreamSynchronize(st));
            if (flags[0] & 2) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_tables: a table is not ordered by (qgene, sgene) (use rcq_graph_pickle_groups)");
            n_edges = (int64_t)(int32_t)(num[0] & 0xFFFFFFFF);
            if (2 * n_edges >= (1ll << 31)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle_tables: 2^30 distinct edges or more");
            // ---- adjacency lists: two entries per edge, stable sort by owning node
            const int64_t M2 = 2 * n_edges;
            uint32_t *d_ek = nullptr, *d_ev = nullptr, *d_ek2 = nullptr, *d_ev2 = nullptr;
            RC(D.get(&d_ek, M2)); RC(D.get(&d_ev, M2)); RC(D.get(&d_ek2, M2)); RC(D.get(&d_ev2, M2));
            LAUNCH(ctx, k_g_entries, grid_for(n_edges, 256), 256, 0, (const uint64_t*)d_edges, n_edges, (uint32_t)n_nodes, d_ek, d_ev, d_flags);
            cub::DoubleBuffer<uint32_t> ek(d_ek, d_ek2), ev(d_ev, d_ev2);
            size_t b5 = 0;
            const int kbits = g_bits_for(n_nodes + 1);
            CU(cub::DeviceRadixSort::SortPairs(nullptr, b5, ek, ev, (int)M2, 0, kbits, st));
            char* ws3 = nullptr;
            RC(D.get(&ws3, (int64_t)b5));
            CU(cub::DeviceRadixSort::SortPairs(ws3, b5, ek, ev, (int)M2, 0, kbits, st));
            ctx->launches += 6;
            RC(to_host(ctx, flags, (const unsigned long long*)d_flags, 4));
            CU(cudaStreamSynchronize(st));
            const int64_t M = M2 - (int64_t)flags[1];
            int32_t* d_nbr = nullptr; int64_t* d_start = nullptr; uint32_t *d_cnt = nullptr, *d_before = nullptr, *d_erel_e = nullptr, *d_erel = nullptr;
            RC(D.get(&d_nbr, M)); RC(D.get(&d_start, n_nodes + 1)); RC(D.get(&d_cnt, n_nodes)); RC(D.get(&d_before, n_nodes));
            RC(D.get(&d_erel_e, n_edges)); RC(D.get(&d_erel, M));
            LAUNCH(ctx, k_g_csr, grid_for(M, 256), 256, 0, (const uint32_t*)ek.Current(), (const uint32_t*)ev.Current(), (const uint64_t*)d_edges, M,
                   n_nodes, d_nbr, d_start);
            LAUNCH(ctx, k_g_first_visits, grid_for(n_nodes, 256), 256, 0, (const int64_t*)d_start, (const int32_t*)d_nbr, n_nodes, d_cnt);
            size_t b6 = 0;
            CU(cub::DeviceScan::ExclusiveSum(nullptr, b6, (const uint32_t*)d_cnt, d_before, (int)n_nodes, st));
            char* ws4 = nullptr;
            RC(D.get(&ws4, (int64_t)b6));
            CU(cub::DeviceScan::ExclusiveSum(ws4, b6, (const uint32_t*)d_cnt, d_before, (int)n_nodes, st));
            LAUNCH(ctx, k_g_edge_memo, grid_for(n_nodes, 256), 256, 0, (const int64_t*)d_start, (const int32_t*)d_nbr, (const uint32_t*)ev.Current(),
                   (const uint32_t*)d_before, n_nodes, d_erel_e);
            LAUNCH(ctx, k_g_entry_memo, grid_for(M, 256), 256, 0, (const uint32_t*)ek.Current(), (const int32_t*)d_nbr, (const uint32_t*)ev.Current(),
                   (const uint32_t*)d_erel_e, M, d_erel);
            if (trace) CU(cudaStreamSynchronize(st));
            lap("device prep");
            // five arrays into the context's page-locked staging (one DMA transfer each; into fresh pageable vectors the
            // same 350 MB took 0.14 s), which the emitter reads in place
            const size_t o_ns = 0, o_ng = o_ns + (size_t)n_nodes * 4, o_st = (o_ng + (size_t)n_nodes * 4 + 7) & ~(size_t)7,
                         o_nb = o_st + (size_t)(n_nodes + 1) * 8, o_er = o_nb + (size_t)M * 4, total = o_er + (size_t)M * 4;
            RC(stage_reserve(ctx, total));
            char* H = ctx->h_stage;
            CU(cudaMemcpyAsync(H + o_ns, d_ns, (size_t)n_nodes * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(H + o_ng, d_ng, (size_t)n_nodes * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(H + o_st, d_start, (size_t)(n_nodes + 1) * 8, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(H + o_nb, d_nbr, (size_t)M * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(H + o_er, d_erel, (size_t)M * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            p_ns = (const int32_t*)(H + o_ns); p_ng = (const int32_t*)(H + o_ng); p_st = (const int64_t*)(H + o_st);
            p_nb = (const int32_t*)(H + o_nb); p_er = (const uint32_t*)(H + o_er);
            lap("to host");
        }
        return rcq_graph_emit(path, n_samples, sample_names, n_nodes, p_ns, p_ng, p_st, p_nb, p_er, n_edges, counts);
    } catch (const std::bad_alloc&) {
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle_tables: out of host memory");
    }
}
