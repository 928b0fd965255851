// phase2.cu — kernel (4): gene-match graph -> connected components (union-find) -> ideal
// components -> per-table masked reductions.  Replaces the Python/networkx/pandas loops of
//   build_graph.py:40-68          (node = (sample, gene), one undirected edge per table row)
//   filtered_distance.py:25-39    (ideal <=> |V| == #samples and 2|E| == |V|(|V|-1))
//   filtered_distance.py:171-210  (sample_count, valid, restricted)
//   similarity_computer.py:21-42  (sum nident, sum length - sum gaps per table)
// Node id = sample*stride + gene (flat arrays).  HBM-bound; algorithmic bytes per table row:
// 8 B (two gene ids) read twice (union, edge count) + 12 B (nident, length, gaps) + two
// validity probes.
#include "search.cuh"

#define EMPTY64 0xFFFFFFFFFFFFFFFFull

struct P2View {
    int32_t n_samples, n_tables;
    int64_t stride, n_rows;
    const int64_t* row_off;           // [n_tables+1]
    const int32_t *t_s, *t_q;         // [n_tables]
    const uint8_t* t_flags;           // [n_tables] or NULL: bit0 rows are graph edges, bit1 rows are summed
    const int32_t *sgene, *qgene, *nident, *length, *gaps;  // [n_rows]
    int32_t* parent;                  // [n_samples*stride]
    uint8_t* present;                 // [n_samples*stride]
    int32_t *vcount, *ecount;         // per root
    uint64_t* ekeys; uint64_t emask;  // unique-edge hash set
};

__device__ __forceinline__ int32_t table_of(const int64_t* __restrict__ row_off, int32_t n_tables, int64_t r) {
    int32_t lo = 0, hi = n_tables;  // row_off[lo] <= r < row_off[hi]
    while (hi - lo > 1) {
        int32_t mid = (lo + hi) >> 1;
        if (__ldg(row_off + mid) <= r) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ int32_t uf_find(int32_t* parent, int32_t x) {
    // path halving; parent pointers only ever decrease towards the root (root = smallest id)
    int32_t p = __ldcg(parent + x);   // L2 reads: other SMs hook/compress concurrently
    while (p != x) {
        int32_t gp = __ldcg(parent + p);
        if (gp != p) parent[x] = gp;
        x = p; p = gp;
    }
    return x;
}
__device__ __forceinline__ void uf_union(int32_t* parent, int32_t a, int32_t b) {
    while (true) {
        a = uf_find(parent, a); b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { int32_t t = a; a = b; b = t; }   // hook the larger root under the smaller
        int32_t old = atomicCAS(&parent[a], a, b);
        if (old == a) return;
    }
}

__global__ void k_p2_init(int32_t* parent, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) parent[i] = (int32_t)i;
}
__global__ void k_p2_union(P2View V) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= V.n_rows) return;
    int32_t t = table_of(V.row_off, V.n_tables, r);
    if (V.t_flags && !(V.t_flags[t] & 1)) return;
    // a row naming the gene pair of the row before it (the forward and the backward row of one reciprocal hit) adds
    // nothing to the graph: half of the atomics of this kernel
    if (r > V.row_off[t] && V.sgene[r] == V.sgene[r - 1] && V.qgene[r] == V.qgene[r - 1]) return;
    int32_t u = (int32_t)(V.t_s[t] * V.stride + V.sgene[r]);
    int32_t v = (int32_t)(V.t_q[t] * V.stride + V.qgene[r]);
    V.present[u] = 1; V.present[v] = 1;
    uf_union(V.parent, u, v);
    // unique undirected edges (nx.Graph collapses repeated rows of the same gene pair)
    uint64_t key = u < v ? ((uint64_t)(uint32_t)u << 32) | (uint32_t)v : ((uint64_t)(uint32_t)v << 32) | (uint32_t)u;
    uint64_t s = (key * 0x9E3779B97F4A7C15ull >> 20) & V.emask;
    while (true) {
        uint64_t cur = V.ekeys[s];
        if (cur == key) break;
        if (cur == EMPTY64) {
            uint64_t old = atomicCAS((unsigned long long*)&V.ekeys[s], EMPTY64, key);
            if (old == EMPTY64 || old == key) break;
        }
        s = (s + 1) & V.emask;
    }
}
__global__ void k_p2_flatten(P2View V, int64_t n_nodes, unsigned long long* counts /* [0]=#comp [3]=#nodes */,
                             int32_t* sample_present) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes || !V.present[i]) return;
    int32_t root = uf_find(V.parent, (int32_t)i);
    V.parent[i] = root;
    atomicAdd(&V.vcount[root], 1);
    if (root == i) atomicAdd(counts, 1ull);
    atomicAdd(counts + 3, 1ull);
    sample_present[i / V.stride] = 1;
}
__global__ void k_p2_edges(P2View V, int64_t cap, unsigned long long* counts /* [4]=#edges */) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= cap) return;
    uint64_t key = V.ekeys[s];
    if (key == EMPTY64) return;
    int32_t u = (int32_t)(key >> 32);
    atomicAdd(&V.ecount[V.parent[u]], 1);   // parent[] is flat after k_p2_flatten
    atomicAdd(counts + 4, 1ull);
}
// valid[node] = node lies in an ideal component
__global__ void k_p2_valid(P2View V, int64_t n_nodes, int32_t sample_count, uint8_t* valid,
                           unsigned long long* counts /* [1]=#ideal [5]=#components with >= sample_count nodes */) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    uint8_t ok = 0;
    if (V.present[i]) {
        int32_t root = V.parent[i];
        int64_t v = V.vcount[root], e = V.ecount[root];
        ok = (v == sample_count) && (2 * e == v * (v - 1));
        if (ok && root == i) atomicAdd(counts + 1, 1ull);
        if (root == i && v >= sample_count) atomicAdd(counts + 5, 1ull);
    }
    valid[i] = ok;
}
// per-table sums over rows whose two genes are both valid; one block-level reduction per table
// segment, int64 accumulation
__global__ void k_p2_reduce(P2View V, const uint8_t* __restrict__ valid, unsigned long long* nident_sum,
                            unsigned long long* denom_sum) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int32_t t = -1;
    unsigned long long a = 0, b = 0;
    if (r < V.n_rows) {
        t = table_of(V.row_off, V.n_tables, r);
        int64_t u = V.t_s[t] * V.stride + V.sgene[r], v = V.t_q[t] * V.stride + V.qgene[r];
        if ((!V.t_flags || (V.t_flags[t] & 2)) && valid[u] && valid[v]) { a = (unsigned long long)V.nident[r]; b = (unsigned long long)(V.length[r] - V.gaps[r]); }
    }
    // warp-level segmented add: rows of a table are contiguous, so the lanes of one table form a
    // contiguous run and a shuffle-down tree restricted to equal table ids is exact
    int lane = threadIdx.x & 31;
    for (int o = 16; o; o >>= 1) {
        unsigned long long a2 = __shfl_down_sync(0xffffffffu, a, o), b2 = __shfl_down_sync(0xffffffffu, b, o);
        int t2 = __shfl_down_sync(0xffffffffu, t, o);
        if (lane + o < 32 && t2 == t) { a += a2; b += b2; }
    }
    int tprev = __shfl_up_sync(0xffffffffu, t, 1);
    bool leader = (lane == 0) || (tprev != t);
    if (t >= 0 && leader && (a | b)) { atomicAdd(&nident_sum[t], a); atomicAdd(&denom_sum[t], b); }
}

// the five columns phase 2 reads, gathered from device-resident tables into one column-major block
__global__ void k_p2_gather(const int64_t* __restrict__ row_off, const int32_t* const* __restrict__ tcols,
                            int32_t n_tables, int64_t n, int32_t* __restrict__ out) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int32_t t = table_of(row_off, n_tables, r);
    const int64_t lo = __ldg(row_off + t), tn = __ldg(row_off + t + 1) - lo;
    const int32_t* c = tcols[t] + (r - lo);
    out[r] = __ldg(c + RCQ_COL_SGENE * tn);
    out[n + r] = __ldg(c + RCQ_COL_QGENE * tn);
    out[2 * n + r] = __ldg(c + RCQ_COL_NIDENT * tn);
    out[3 * n + r] = __ldg(c + RCQ_COL_LENGTH * tn);
    out[4 * n + r] = __ldg(c + RCQ_COL_GAPS * tn);
}

// ---------------------------------------------------------------------------------------------
static int phase2_run(rcq_ctx* ctx, P2View V, int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum,
                      uint8_t* valid_out, int64_t counts[6], const uint8_t* h_flags = nullptr) {
    const int64_t n_nodes = (int64_t)V.n_samples * V.stride;
    if (n_nodes >= (1ll << 31)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_phase2: n_samples*stride exceeds 2^31 nodes");
    unsigned long long* ctr = (unsigned long long*)ctx->d_ctr;
    int64_t ecap = 1024;
    while (ecap < 2 * V.n_rows) ecap <<= 1;
    V.emask = (uint64_t)(ecap - 1);
    uint8_t* valid = nullptr; int32_t* spresent = nullptr;
    unsigned long long *d_ni = nullptr, *d_de = nullptr;
    RC(dalloc(ctx, &V.parent, n_nodes));
    RC(dalloc(ctx, &V.present, n_nodes));
    RC(dalloc(ctx, &V.vcount, n_nodes));
    RC(dalloc(ctx, &V.ecount, n_nodes));
    RC(dalloc(ctx, &V.ekeys, ecap));
    RC(dalloc(ctx, &valid, n_nodes));
    RC(dalloc(ctx, &spresent, V.n_samples));
    RC(dalloc(ctx, &d_ni, V.n_tables));
    RC(dalloc(ctx, &d_de, V.n_tables));
    CU(cudaMemsetAsync(V.present, 0, n_nodes, ctx->stream));
    CU(cudaMemsetAsync(V.vcount, 0, n_nodes * 4, ctx->stream));
    CU(cudaMemsetAsync(V.ecount, 0, n_nodes * 4, ctx->stream));
    CU(cudaMemsetAsync(V.ekeys, 0xFF, ecap * 8, ctx->stream));
    CU(cudaMemsetAsync(spresent, 0, (size_t)std::max(V.n_samples, 1) * 4, ctx->stream));
    CU(cudaMemsetAsync(d_ni, 0, (size_t)std::max(V.n_tables, 1) * 8, ctx->stream));
    CU(cudaMemsetAsync(d_de, 0, (size_t)std::max(V.n_tables, 1) * 8, ctx->stream));
    CU(cudaMemsetAsync(ctr, 0, 16 * sizeof(int64_t), ctx->stream));
    stage_tic(ctx, 6);
    LAUNCH(ctx, k_p2_init, grid_for(n_nodes, 256), 256, 0, V.parent, n_nodes);
    if (V.n_rows) LAUNCH(ctx, k_p2_union, grid_for(V.n_rows, 256), 256, 0, V);
    LAUNCH(ctx, k_p2_flatten, grid_for(n_nodes, 256), 256, 0, V, n_nodes, ctr, spresent);
    LAUNCH(ctx, k_p2_edges, grid_for(ecap, 256), 256, 0, V, ecap, ctr);
    if (sample_count <= 0) {
        std::vector<int32_t> sp(std::max(V.n_samples, 1));
        CU(cudaMemcpyAsync(sp.data(), spresent, (size_t)V.n_samples * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        sample_count = 0;
        for (int i = 0; i < V.n_samples; i++) sample_count += sp[i] != 0;
    }
    LAUNCH(ctx, k_p2_valid, grid_for(n_nodes, 256), 256, 0, V, n_nodes, sample_count, valid, ctr);
    if (V.n_rows) LAUNCH(ctx, k_p2_reduce, grid_for(V.n_rows, 256), 256, 0, V, valid, d_ni, d_de);
    stage_toc(ctx, 6);
    if (V.n_tables) {
        CU(cudaMemcpyAsync(nident_sum, d_ni, (size_t)V.n_tables * 8, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaMemcpyAsync(denom_sum, d_de, (size_t)V.n_tables * 8, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (valid_out) CU(cudaMemcpyAsync(valid_out, valid, n_nodes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_pin, ctr, 8 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    stage_collect(ctx);
    if (counts) {
        counts[0] = ctx->h_pin[0]; counts[1] = ctx->h_pin[1]; counts[2] = sample_count;
        counts[3] = ctx->h_pin[3]; counts[4] = ctx->h_pin[4]; counts[5] = ctx->h_pin[5];
    }
    dfree(ctx, V.parent); dfree(ctx, V.present); dfree(ctx, V.vcount); dfree(ctx, V.ecount);
    dfree(ctx, V.ekeys); dfree(ctx, valid); dfree(ctx, spresent); dfree(ctx, d_ni); dfree(ctx, d_de);
    for (int32_t t = 0; t < V.n_tables; t++)
        if (denom_sum[t] == 0 && (!h_flags || (h_flags[t] & 2))) return rcq_fail(RCQ_ERR_NO_IDEAL, "rcq_phase2: a table has no row inside ideal components");
    return 0;
}

extern "C" int rcq_phase2_host(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables,
                               const int64_t* row_off, const int32_t* t_s, const int32_t* t_q,
                               const int32_t* sgene, const int32_t* qgene, const int32_t* nident,
                               const int32_t* length, const int32_t* gaps, const uint8_t* t_flags,
                               int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum, uint8_t* valid,
                               int64_t counts[6]) {
    if (!ctx || n_samples < 0 || stride < 1 || n_tables < 0 || (n_tables && (!row_off || !t_s || !t_q || !nident_sum || !denom_sum)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_phase2_host: bad argument");
    CU(cudaSetDevice(ctx->device));
    P2View V;
    memset(&V, 0, sizeof V);
    V.n_samples = n_samples; V.n_tables = n_tables; V.stride = stride;
    V.n_rows = n_tables ? row_off[n_tables] : 0;
    const int64_t n = V.n_rows;
    int64_t* d_off = nullptr; int32_t *d_ts = nullptr, *d_tq = nullptr, *d_cols = nullptr;
    RC(dalloc(ctx, &d_off, n_tables + 1));
    RC(dalloc(ctx, &d_ts, n_tables));
    RC(dalloc(ctx, &d_tq, n_tables));
    RC(dalloc(ctx, &d_cols, 5 * n));
    uint8_t* d_flags = nullptr;
    if (t_flags && n_tables) {
        RC(dalloc(ctx, &d_flags, n_tables));
        CU(cudaMemcpyAsync(d_flags, t_flags, (size_t)n_tables, cudaMemcpyHostToDevice, ctx->stream));
    }
    V.t_flags = d_flags;
    if (n_tables) {
        CU(cudaMemcpyAsync(d_off, row_off, (size_t)(n_tables + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(d_ts, t_s, (size_t)n_tables * 4, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(d_tq, t_q, (size_t)n_tables * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    const int32_t* src[5] = {sgene, qgene, nident, length, gaps};
    for (int c = 0; c < 5 && n; c++)
        CU(cudaMemcpyAsync(d_cols + c * n, src[c], (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    V.row_off = d_off; V.t_s = d_ts; V.t_q = d_tq;
    V.sgene = d_cols; V.qgene = d_cols + n; V.nident = d_cols + 2 * n; V.length = d_cols + 3 * n; V.gaps = d_cols + 4 * n;
    int rc = phase2_run(ctx, V, sample_count, nident_sum, denom_sum, valid, counts, d_flags ? t_flags : nullptr);
    dfree(ctx, d_off); dfree(ctx, d_ts); dfree(ctx, d_tq); dfree(ctx, d_cols); dfree(ctx, d_flags);
    return rc;
}

extern "C" int rcq_phase2_tables(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables,
                                 const rcq_table* const* tables, const int32_t* t_s, const int32_t* t_q,
                                 int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum,
                                 uint8_t* valid, int64_t counts[6]) {
    if (!ctx || n_samples < 0 || stride < 1 || n_tables < 0 || (n_tables && (!tables || !t_s || !t_q || !nident_sum || !denom_sum)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_phase2_tables: bad argument");
    CU(cudaSetDevice(ctx->device));
    // one host block [row offsets | table base pointers | t_s | t_q] -> one copy; the five columns the pass
    // reads are then gathered from the tables by one kernel (not 5 copies per table)
    const size_t words64 = (size_t)(n_tables + 1) + (size_t)n_tables;
    std::vector<int64_t> host(words64 + (size_t)n_tables + 1, 0);      // + n_tables int32 pairs
    int64_t* off = host.data();
    for (int t = 0; t < n_tables; t++) off[t + 1] = off[t] + tables[t]->n;
    for (int t = 0; t < n_tables; t++) host[n_tables + 1 + t] = (int64_t)(uintptr_t)tables[t]->cols;
    int32_t* h32 = (int32_t*)(host.data() + words64);
    for (int t = 0; t < n_tables; t++) { h32[t] = t_s[t]; h32[n_tables + t] = t_q[t]; }
    const int64_t n = off[n_tables];
    P2View V;
    memset(&V, 0, sizeof V);
    V.n_samples = n_samples; V.n_tables = n_tables; V.stride = stride; V.n_rows = n;
    int64_t* d_host = nullptr; int32_t* d_cols = nullptr;
    RC(dalloc(ctx, &d_host, (int64_t)host.size()));
    RC(dalloc(ctx, &d_cols, 5 * n));
    CU(cudaMemcpyAsync(d_host, host.data(), host.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    const int64_t* d_off = d_host;
    const int32_t* d_ts = (const int32_t*)(d_host + words64);
    const int32_t* d_tq = d_ts + n_tables;
    if (n) LAUNCH(ctx, k_p2_gather, grid_for(n, 256), 256, 0, d_off, (const int32_t* const*)(d_host + n_tables + 1), n_tables, n, d_cols);
    V.row_off = d_off; V.t_s = d_ts; V.t_q = d_tq;
    V.sgene = d_cols; V.qgene = d_cols + n; V.nident = d_cols + 2 * n; V.length = d_cols + 3 * n; V.gaps = d_cols + 4 * n;
    int rc = phase2_run(ctx, V, sample_count, nident_sum, denom_sum, valid, counts);
    dfree(ctx, d_host); dfree(ctx, d_cols);
    return rc;
}

// ---- multi-GPU exchange of what phase 2 reads: five columns, 20 bytes per row -------------
// rcq_phase2_pack: the five columns of this rank's tables as one column-major block [5][n_local] (k_p2_gather);
// the host layer all-gathers the blocks over NCCL.  rcq_phase2_packed: phase 2 on the gathered buffer; table t
// has its rows at dev_packed[src_base[t] + c * src_stride[t] + x] for column c, row x (src_stride = rows of the
// block that holds it).
__global__ void k_p2_unpack(const int64_t* __restrict__ row_off, const int64_t* __restrict__ src_base,
                            const int64_t* __restrict__ src_stride, const int32_t* __restrict__ in, int32_t n_tables,
                            int64_t n, int32_t* __restrict__ out) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int32_t t = table_of(row_off, n_tables, r);
    const int64_t at = __ldg(src_base + t) + (r - __ldg(row_off + t)), st = __ldg(src_stride + t);
#pragma unroll
    for (int c = 0; c < 5; c++) out[c * n + r] = __ldg(in + at + c * st);
}

extern "C" int rcq_phase2_pack(rcq_ctx* ctx, int32_t n_tables, const rcq_table* const* tables, int32_t* dev_out) {
    if (!ctx || n_tables < 0 || (n_tables && (!tables || !dev_out))) return rcq_fail(RCQ_ERR_ARG, "rcq_phase2_pack: bad argument");
    CU(cudaSetDevice(ctx->device));
    std::vector<int64_t> host((size_t)2 * n_tables + 1, 0);
    for (int t = 0; t < n_tables; t++) host[t + 1] = host[t] + tables[t]->n;
    for (int t = 0; t < n_tables; t++) host[n_tables + 1 + t] = (int64_t)(uintptr_t)tables[t]->cols;
    const int64_t n = host[n_tables];
    if (n) {
        int64_t* d_host = nullptr;
        RC(dalloc(ctx, &d_host, (int64_t)host.size()));
        CU(cudaMemcpyAsync(d_host, host.data(), host.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
        LAUNCH(ctx, k_p2_gather, grid_for(n, 256), 256, 0, d_host, (const int32_t* const*)(d_host + n_tables + 1), n_tables, n, dev_out);
        dfree(ctx, d_host);
    }
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int rcq_phase2_packed(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables, const int64_t* row_off,
                                 const int32_t* t_s, const int32_t* t_q, const int64_t* src_base, const int64_t* src_stride,
                                 const int32_t* dev_packed, int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum,
                                 uint8_t* valid, int64_t counts[6]) {
    if (!ctx || n_samples < 0 || stride < 1 || n_tables < 0 ||
        (n_tables && (!row_off || !t_s || !t_q || !src_base || !src_stride || !nident_sum || !denom_sum)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_phase2_packed: bad argument");
    CU(cudaSetDevice(ctx->device));
    const int64_t n = n_tables ? row_off[n_tables] : 0;
    if (n && !dev_packed) return rcq_fail(RCQ_ERR_ARG, "rcq_phase2_packed: bad argument");
    // one host block [row offsets | src_base | src_stride | t_s | t_q]
    const size_t words64 = (size_t)3 * n_tables + 1;
    std::vector<int64_t> host(words64 + (size_t)n_tables + 1, 0);
    for (int t = 0; t <= n_tables && n_tables; t++) host[t] = row_off[t];
    for (int t = 0; t < n_tables; t++) { host[n_tables + 1 + t] = src_base[t]; host[2 * (size_t)n_tables + 1 + t] = src_stride[t]; }
    int32_t* h32 = (int32_t*)(host.data() + words64);
    for (int t = 0; t < n_tables; t++) { h32[t] = t_s[t]; h32[n_tables + t] = t_q[t]; }
    P2View V;
    memset(&V, 0, sizeof V);
    V.n_samples = n_samples; V.n_tables = n_tables; V.stride = stride; V.n_rows = n;
    int64_t* d_host = nullptr; int32_t* d_cols = nullptr;
    RC(dalloc(ctx, &d_host, (int64_t)host.size()));
    RC(dalloc(ctx, &d_cols, 5 * n));
    CU(cudaMemcpyAsync(d_host, host.data(), host.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (n) LAUNCH(ctx, k_p2_unpack, grid_for(n, 256), 256, 0, d_host, d_host + n_tables + 1, d_host + 2 * n_tables + 1, dev_packed, n_tables, n, d_cols);
    V.row_off = d_host; V.t_s = (const int32_t*)(d_host + words64); V.t_q = V.t_s + n_tables;
    V.sgene = d_cols; V.qgene = d_cols + n; V.nident = d_cols + 2 * n; V.length = d_cols + 3 * n; V.gaps = d_cols + 4 * n;
    int rc = phase2_run(ctx, V, sample_count, nident_sum, denom_sum, valid, counts);
    dfree(ctx, d_host); dfree(ctx, d_cols);
    return rc;
}

// ---- flat (de)serialisation for the NCCL all-gather of hit tables ----------------
extern "C" int64_t rcq_table_flat_words(const rcq_table* t) { return t ? (int64_t)RCQ_NCOLS * t->n : 0; }
extern "C" int rcq_table_flatten(rcq_ctx* ctx, const rcq_table* t, int32_t* dev_out) {
    if (!ctx || !t) return rcq_fail(RCQ_ERR_ARG, "rcq_table_flatten: bad argument");
    CU(cudaSetDevice(ctx->device));
    if (t->n) CU(cudaMemcpyAsync(dev_out, t->cols, (size_t)RCQ_NCOLS * t->n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int rcq_table_from_flat(rcq_ctx* ctx, const int32_t* dev_in, int64_t n_rows, rcq_table** out) {
    if (!ctx || !out || n_rows < 0) return rcq_fail(RCQ_ERR_ARG, "rcq_table_from_flat: bad argument");
    CU(cudaSetDevice(ctx->device));
    rcq_table* T = new rcq_table();
    T->ctx = ctx; T->n = n_rows;
    if (n_rows) {
        RC(dalloc(ctx, &T->cols, (int64_t)RCQ_NCOLS * n_rows));
        CU(cudaMemcpyAsync(T->cols, dev_in, (size_t)RCQ_NCOLS * n_rows * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    *out = T;
    return 0;
}
