// host_graph.cu — graph.pkl written straight from the flat edge arrays (host only, no device work).
// The reference pickles the networkx graph of build_graph (build_graph.py:40-68) with
// pickle.dump(graph, f, HIGHEST_PROTOCOL) (filtering_step.py:158-159); building that dict-of-dict
// object in Python and pickling it costs seconds per million edges.  This writer emits the pickle
// stream of the SAME object directly: `networkx.classes.graph.Graph()` + a state dict with
//   graph = {}, _node = {node: {}}, _adj = {node: {neighbour: attr}}, __networkx_cache__ = {}
// where node = (sample path string, gene id int), nodes appear in first-appearance order (u before v,
// row by row), neighbours in edge-insertion order, repeated rows collapse, and adj[u][v] and adj[v][u]
// are ONE shared (memoised) empty attribute dict — exactly what add_edges_from builds.
// Opcodes are pickle protocol 4 (readable by every Python 3.4+): only strings, node tuples and edge
// dicts are memoised, because nothing else is referenced twice.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include "common.cuh"
#include "host_graph.h"

namespace {
// Append-only byte buffer with bulk capacity checks: the adjacency section is hundreds of megabytes, one
// bounds-checked push_back per byte would dominate.
struct PickleOut {
    std::vector<uint8_t> b;
    size_t n = 0;
    void need(size_t extra) { if (n + extra > b.size()) b.resize(std::max(b.size() * 2, n + extra + (1u << 16))); }
    void op(uint8_t c) { b[n++] = c; }
    void raw(const void* p, size_t k) { memcpy(b.data() + n, p, k); n += k; }
    void u32(uint32_t v) { memcpy(b.data() + n, &v, 4); n += 4; }      // little-endian host
    void memoize() { op(0x94); }                                           // MEMOIZE
    void get(uint32_t id) { if (id < 256) { op('h'); op((uint8_t)id); } else { op('j'); u32(id); } }   // BINGET / LONG_BINGET
    void str(const char* s, size_t k) {                                    // SHORT_BINUNICODE / BINUNICODE (utf-8 bytes)
        need(k + 8);
        if (k < 256) { op(0x8c); op((uint8_t)k); } else { op('X'); u32((uint32_t)k); }
        raw(s, k);
    }
    void integer(int32_t v) {
        if (v >= 0 && v < 256) { op('K'); op((uint8_t)v); }
        else if (v >= 0 && v < 65536) { op('M'); op((uint8_t)v); op((uint8_t)(v >> 8)); }
        else { op('J'); u32((uint32_t)v); }
    }
    void empty_dict() { op('}'); }
    void mark() { op('('); }
    void setitems() { op('u'); }
};

template <typename F>
void parallel_for(int64_t n, int n_threads, F f) {      // f(part, begin, end)
    if (n_threads <= 1 || n < 4096) { f(0, (int64_t)0, n); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++)
        th.emplace_back([=] { f(t, n * t / n_threads, n * (t + 1) / n_threads); });
    for (auto& x : th) x.join();
}
template <typename F>
void run_threads(int n_threads, F f) {                  // f(t) on n_threads threads
    if (n_threads <= 1) { f(0); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++) th.emplace_back([=] { f(t); });
    for (auto& x : th) x.join();
}
}  // namespace

// ---- the stream.  Shared by this file's host front end and the device front end (graph_device.cu).  Nodes in
// first-appearance order; adjacency lists (CSR start/nbr) in edge-insertion order; erel[a] = RCQ_GRAPH_FIRST_VISIT
// when entry a is the first visit of its edge (nbr[a] >= the node that owns the list: the attribute dict is created
// and memoised there), else the memo number of that dict relative to the end of the _node section.
int rcq_graph_emit(const char* path, int32_t n_samples, const char* const* sample_names, int64_t n_nodes,
                   const int32_t* node_sample, const int32_t* node_gene, const int64_t* start, const int32_t* nbr,
                   const uint32_t* erel, int64_t n_edges, int64_t counts[3]) {
    const int n_threads = (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    static const bool trace = getenv("RCQ_TRACE") != nullptr;
    auto T0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        auto T1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[rcq_graph_pickle] %-14s %.3f s\n", what, std::chrono::duration<double>(T1 - T0).count());
        T0 = T1;
    };
    if (n_nodes + n_edges + n_samples + 16 >= (1ll << 32)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle: more than 2^32 memo entries");
    // strings and node tuples are memoised in the _node section, in node order
    std::vector<uint32_t> sample_memo((size_t)n_samples, 0xFFFFFFFFu), node_memo((size_t)n_nodes);
    uint32_t memo = 0;
    for (int64_t i = 0; i < n_nodes; i++) {
        if (sample_memo[node_sample[i]] == 0xFFFFFFFFu) sample_memo[node_sample[i]] = memo++;
        node_memo[i] = memo++;
    }
    // ---- the stream: head (+ _node section) sequentially, the _adj section in parallel parts, tail
    PickleOut H;
    H.need((size_t)(256 + 24 * n_nodes));
    H.op(0x80); H.op(4);                                                    // PROTO 4
    H.str("networkx.classes.graph", 22); H.str("Graph", 5); H.op(0x93);     // STACK_GLOBAL
    H.op(')'); H.op('R');                                                   // EMPTY_TUPLE REDUCE -> Graph()
    H.empty_dict(); H.mark();                                               // state = {
    H.str("graph", 5); H.empty_dict();
    H.str("_node", 5); H.empty_dict(); H.mark();
    {
        std::vector<uint8_t> seen((size_t)n_samples, 0);
        for (int64_t i = 0; i < n_nodes; i++) {
            const int32_t s = node_sample[i];
            H.need(32);
            if (!seen[s]) { H.str(sample_names[s], strlen(sample_names[s])); H.need(32); H.memoize(); seen[s] = 1; }
            else H.get(sample_memo[s]);
            H.integer(node_gene[i]);
            H.op(0x86);                                                     // TUPLE2
            H.memoize();
            H.empty_dict();
        }
    }
    lap("node section");
    H.need(32);
    H.setitems();
    H.str("_adj", 4); H.need(8); H.empty_dict(); H.mark();
    std::vector<PickleOut> parts((size_t)n_threads);
    int used_parts = 1;
    if (n_nodes >= 4096) used_parts = n_threads;
    parallel_for(n_nodes, used_parts, [&](int t, int64_t i0, int64_t i1) {
        PickleOut& P = parts[t];
        P.need((size_t)(16 * (i1 - i0) + 10 * (start[i1] - start[i0]) + 64));
        for (int64_t i = i0; i < i1; i++) {
            P.need((size_t)(16 + 10 * (start[i + 1] - start[i])));
            P.get(node_memo[i]);
            P.empty_dict();
            if (start[i + 1] > start[i]) {
                P.mark();
                for (int64_t a = start[i]; a < start[i + 1]; a++) {
                    P.get(node_memo[nbr[a]]);
                    if (erel[a] == RCQ_GRAPH_FIRST_VISIT) { P.empty_dict(); P.memoize(); } else P.get(memo + erel[a]);
                }
                P.setitems();
            }
        }
    });
    lap("stream");
    PickleOut T;
    T.need(64);
    T.setitems();
    T.str("__networkx_cache__", 18); T.need(8); T.empty_dict();
    T.setitems();                                                           // }
    T.op('b'); T.op('.');                                                   // BUILD STOP
    FILE* f = fopen(path, "wb");
    if (!f) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: cannot open the output file");
    size_t total = 0, wrote = 0;
    auto put = [&](const PickleOut& P) { total += P.n; if (P.n) wrote += fwrite(P.b.data(), 1, P.n, f); };
    put(H);
    for (int t = 0; t < used_parts; t++) put(parts[t]);
    put(T);
    const bool closed = fclose(f) == 0;
    lap("file write");
    if (!closed || wrote != total) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: short write");
    if (counts) { counts[0] = n_nodes; counts[1] = n_edges; counts[2] = (int64_t)total; }
    return 0;
}


// Groups g = 0..n_groups-1 hold rows [g_off[g], g_off[g+1]) of (sgene, qgene) between samples g_s[g] and
// g_q[g] (indices into sample_names); isolated nodes (degree 0) follow the edges.  counts[0] = nodes,
// counts[1] = distinct undirected edges, counts[2] = bytes written.
//
// Cost at BASELINE configs[2] (435 tables, 43.5 million rows, 21.75 million edges, a 0.4 GB stream): node ids in
// one sequential pass (first-appearance order is part of the contract), everything else on all host cores -
// duplicate rows are dropped per group (a table sorted by gene pair needs one comparison per row; any other
// group goes through a hash set of its own), adjacency lists are filled per node range, memo numbers of the edge
// dictionaries come from a prefix sum over "first visits", and the adjacency section is written in parallel
// parts.  (The first version - one global hash set, one byte at a time - took 4.6 s of a 6.7 s rna_clique().)
static int graph_pickle_impl(const char* path, int32_t n_samples, const char* const* sample_names,
                             int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_off,
                             const int32_t* const* sgene_of, const int32_t* const* qgene_of, int64_t n_isolated,
                             const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]) {
    // sgene_of[g] / qgene_of[g]: the columns of group g, biased by -g_off[g] so that they are indexed by the global row
    if (!path || n_samples < 0 || n_groups < 0 || n_isolated < 0 || (n_samples && !sample_names) ||
        (n_groups && (!g_s || !g_q || !g_off)) || (n_isolated && (!iso_sample || !iso_gene)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: bad argument");
    const int64_t n_rows = n_groups ? g_off[n_groups] : 0;
    if (n_rows && (!sgene_of || !qgene_of)) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: bad argument");
    const int n_threads = (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    static const bool trace = getenv("RCQ_TRACE") != nullptr;
    auto T0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        auto T1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[rcq_graph_pickle] %-14s %.3f s\n", what, std::chrono::duration<double>(T1 - T0).count());
        T0 = T1;
    };
    // ---- node ids in first-appearance order
    std::vector<int32_t> max_gene((size_t)n_samples, -1);
    for (int32_t g = 0; g < n_groups; g++) {
        if (g_s[g] < 0 || g_s[g] >= n_samples || g_q[g] < 0 || g_q[g] >= n_samples)
            return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: sample index out of range");
        int32_t ms = max_gene[g_s[g]], mq = max_gene[g_q[g]];
        const int32_t* sgene = sgene_of[g]; const int32_t* qgene = qgene_of[g];
        for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) {
            if (sgene[r] < 0 || qgene[r] < 0) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: negative gene id");
            if (sgene[r] > ms) ms = sgene[r];
            if (qgene[r] > mq) mq = qgene[r];
        }
        max_gene[g_s[g]] = std::max(max_gene[g_s[g]], ms);
        max_gene[g_q[g]] = std::max(max_gene[g_q[g]], mq);
    }
    for (int64_t i = 0; i < n_isolated; i++) {
        if (iso_sample[i] < 0 || iso_sample[i] >= n_samples || iso_gene[i] < 0)
            return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: bad isolated node");
        if (iso_gene[i] > max_gene[iso_sample[i]]) max_gene[iso_sample[i]] = iso_gene[i];
    }
    // (sample, gene) -> node id: dense tables when the gene ids are compact (rnaSPAdes numbering), a hash map
    // when a custom id parser produces large sparse ids
    int64_t span = 0;
    for (int32_t s = 0; s < n_samples; s++) span += (int64_t)max_gene[s] + 1;
    const bool dense = span <= 16 * (n_rows + n_isolated) + (1 << 20);
    std::vector<std::vector<int32_t>> id((size_t)n_samples);
    std::unordered_map<uint64_t, int32_t> sparse;
    if (dense) for (int32_t s = 0; s < n_samples; s++) id[s].assign((size_t)(max_gene[s] + 1), -1);
    else sparse.reserve((size_t)(2 * n_rows + n_isolated));
    std::vector<int32_t> node_sample, node_gene;
    auto node_of = [&](int32_t s, int32_t gene) {
        if (dense) {
            int32_t& x = id[s][gene];
            if (x < 0) { x = (int32_t)node_sample.size(); node_sample.push_back(s); node_gene.push_back(gene); }
            return x;
        }
        auto it = sparse.find(((uint64_t)(uint32_t)s << 32) | (uint32_t)gene);
        if (it != sparse.end()) return it->second;
        const int32_t x = (int32_t)node_sample.size();
        sparse.emplace(((uint64_t)(uint32_t)s << 32) | (uint32_t)gene, x);
        node_sample.push_back(s); node_gene.push_back(gene);
        return x;
    };
    std::vector<int32_t> ru((size_t)n_rows), rv((size_t)n_rows);         // node ids of every row
    if (dense && n_rows >= 65536 && 2 * (n_rows + n_isolated) < 0xFFFFFFF0ll) {
        // first appearance of every (sample, gene) as the smallest "row << 1 | side" that names it, found by all
        // threads with an atomic minimum; nodes are then numbered in that order (u before v, row by row)
        std::vector<int64_t> base((size_t)n_samples + 1, 0);
        for (int32_t sidx = 0; sidx < n_samples; sidx++) base[sidx + 1] = base[sidx] + max_gene[sidx] + 1;
        std::vector<std::atomic<uint32_t>> first((size_t)base[n_samples]);
        parallel_for((int64_t)first.size(), n_threads, [&](int, int64_t a0, int64_t a1) {
            for (int64_t a = a0; a < a1; a++) first[a].store(0xFFFFFFFFu, std::memory_order_relaxed);
        });
        auto note = [&](int64_t slot, uint32_t key) {
            uint32_t cur = first[slot].load(std::memory_order_relaxed);
            while (key < cur && !first[slot].compare_exchange_weak(cur, key, std::memory_order_relaxed)) {}
        };
        parallel_for(n_groups, n_threads, [&](int, int64_t g0, int64_t g1) {
            for (int64_t g = g0; g < g1; g++) {
                const int64_t bs = base[g_s[g]], bq = base[g_q[g]];
                const int32_t* sgene = sgene_of[g]; const int32_t* qgene = qgene_of[g];
                for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) {
                    note(bs + sgene[r], (uint32_t)(2 * r));
                    note(bq + qgene[r], (uint32_t)(2 * r + 1));
                }
            }
        });
        for (int64_t i = 0; i < n_isolated; i++) note(base[iso_sample[i]] + iso_gene[i], (uint32_t)(2 * (n_rows + i)));
        std::vector<uint64_t> order;                                     // first-appearance key << 32 | slot
        order.reserve((size_t)std::min<int64_t>(base[n_samples], 2 * n_rows + n_isolated));
        for (int64_t a = 0; a < base[n_samples]; a++) {
            const uint32_t k = first[a].load(std::memory_order_relaxed);
            if (k != 0xFFFFFFFFu) order.push_back(((uint64_t)k << 32) | (uint64_t)a);
        }
        std::sort(order.begin(), order.end());
        node_sample.resize(order.size()); node_gene.resize(order.size());
        std::vector<int32_t> slot_id((size_t)base[n_samples], -1);
        parallel_for((int64_t)order.size(), n_threads, [&](int, int64_t a0, int64_t a1) {
            for (int64_t x = a0; x < a1; x++) {
                const int64_t a = (int64_t)(order[x] & 0xFFFFFFFFull);
                const int32_t sidx = (int32_t)(std::upper_bound(base.begin(), base.end(), a) - base.begin()) - 1;
                node_sample[x] = sidx; node_gene[x] = (int32_t)(a - base[sidx]);
                slot_id[a] = (int32_t)x;
            }
        });
        parallel_for(n_groups, n_threads, [&](int, int64_t g0, int64_t g1) {
            for (int64_t g = g0; g < g1; g++) {
                const int64_t bs = base[g_s[g]], bq = base[g_q[g]];
                const int32_t* sgene = sgene_of[g]; const int32_t* qgene = qgene_of[g];
                for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) { ru[r] = slot_id[bs + sgene[r]]; rv[r] = slot_id[bq + qgene[r]]; }
            }
        });
    } else {
        for (int32_t g = 0; g < n_groups; g++) {
            const int32_t sidx = g_s[g], q = g_q[g];
            const int32_t* sgene = sgene_of[g]; const int32_t* qgene = qgene_of[g];
            for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) { ru[r] = node_of(sidx, sgene[r]); rv[r] = node_of(q, qgene[r]); }
        }
        for (int64_t i = 0; i < n_isolated; i++) node_of(iso_sample[i], iso_gene[i]);
    }
    const int64_t n_nodes = (int64_t)node_sample.size();
    lap("node ids");
    // ---- distinct undirected edges in insertion order.  keep[r] = row r is the first row of its edge.
    // Two groups can only share an edge when they join the same two samples; then one global set decides.
    bool pairs_distinct = true;
    {
        std::vector<uint64_t> pk((size_t)n_groups);
        for (int32_t g = 0; g < n_groups; g++) {
            const uint32_t a = (uint32_t)std::min(g_s[g], g_q[g]), b = (uint32_t)std::max(g_s[g], g_q[g]);
            pk[g] = ((uint64_t)a << 32) | b;
        }
        std::sort(pk.begin(), pk.end());
        pairs_distinct = std::adjacent_find(pk.begin(), pk.end()) == pk.end();
    }
    std::vector<uint8_t> keep((size_t)n_rows, 1);
    auto edge_key = [&](int64_t r) {
        const uint32_t u = (uint32_t)ru[r], v = (uint32_t)rv[r];
        return u < v ? ((uint64_t)u << 32) | v : ((uint64_t)v << 32) | u;
    };
    auto dedup_rows = [&](int64_t r0, int64_t r1) {        // hash set over one range of rows
        size_t cap = 64;
        while (cap < 2 * (size_t)(r1 - r0) + 2) cap <<= 1;
        std::vector<uint64_t> set(cap, ~0ull);
        for (int64_t r = r0; r < r1; r++) {
            const uint64_t key = edge_key(r);
            size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & (cap - 1);
            bool seen = false;
            while (set[h] != ~0ull) { if (set[h] == key) { seen = true; break; } h = (h + 1) & (cap - 1); }
            if (seen) keep[r] = 0; else set[h] = key;
        }
    };
    if (!pairs_distinct) dedup_rows(0, n_rows);
    else
        parallel_for(n_groups, n_threads, [&](int, int64_t g0, int64_t g1) {
            for (int64_t g = g0; g < g1; g++) {
                const int64_t r0 = g_off[g], r1 = g_off[g + 1];
                const int32_t* sgene = sgene_of[g]; const int32_t* qgene = qgene_of[g];
                bool sorted = true;        // rows ordered by (qgene, sgene), as rcq_search_pair and the reference write them
                for (int64_t r = r0 + 1; r < r1 && sorted; r++)
                    sorted = qgene[r] > qgene[r - 1] || (qgene[r] == qgene[r - 1] && sgene[r] >= sgene[r - 1]);
                if (sorted) { for (int64_t r = r0 + 1; r < r1; r++) if (qgene[r] == qgene[r - 1] && sgene[r] == sgene[r - 1]) keep[r] = 0; }
                else dedup_rows(r0, r1);
            }
        });
    // compaction: kept rows of every chunk counted, then copied, both in parallel
    std::vector<int32_t> eu, ev;
    {
        const int ct = n_rows >= 65536 ? n_threads : 1;
        std::vector<int64_t> kept((size_t)ct + 1, 0);
        run_threads(ct, [&](int t) {
            int64_t c = 0;
            for (int64_t r = n_rows * t / ct; r < n_rows * (t + 1) / ct; r++) c += keep[r];
            kept[t + 1] = c;
        });
        for (int t = 0; t < ct; t++) kept[t + 1] += kept[t];
        eu.resize((size_t)kept[ct]); ev.resize((size_t)kept[ct]);
        run_threads(ct, [&](int t) {
            int64_t w = kept[t];
            for (int64_t r = n_rows * t / ct; r < n_rows * (t + 1) / ct; r++) if (keep[r]) { eu[w] = ru[r]; ev[w] = rv[r]; w++; }
        });
    }
    std::vector<int32_t>().swap(ru); std::vector<int32_t>().swap(rv); std::vector<uint8_t>().swap(keep);
    const int64_t n_edges = (int64_t)eu.size();
    lap("distinct edges");
    if (n_nodes + n_edges + n_samples + 16 >= (1ll << 32)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle: more than 2^32 memo entries");
    // ---- adjacency lists in edge-insertion order (CSR; a self-loop is listed once).  Every thread reads all edges in
    // order and files the end points it owns (node id modulo the thread count): lists stay in insertion order,
    // and the scattered writes - the expensive part - are shared out.
    std::vector<int64_t> start((size_t)n_nodes + 1, 0);
    int adj_threads = 1;                                    // a power of two: ownership by the low bits of the node id
    while (n_edges >= 65536 && adj_threads * 2 <= n_threads) adj_threads *= 2;
    const int32_t adj_mask = adj_threads - 1;
    run_threads(adj_threads, [&](int t) {
        for (int64_t e = 0; e < n_edges; e++) {
            const int32_t u = eu[e], v = ev[e];
            if ((u & adj_mask) == t) start[(size_t)u + 1]++;
            if (v != u && (v & adj_mask) == t) start[(size_t)v + 1]++;
        }
    });
    for (int64_t i = 0; i < n_nodes; i++) start[i + 1] += start[i];
    std::vector<int64_t> fill(start.begin(), start.end() - 1);
    std::vector<int32_t> nbr((size_t)start[n_nodes]);
    std::vector<int32_t> via((size_t)start[n_nodes]);
    run_threads(adj_threads, [&](int t) {
        for (int64_t e = 0; e < n_edges; e++) {
            const int32_t u = eu[e], v = ev[e];
            if ((u & adj_mask) == t) { nbr[fill[u]] = v; via[fill[u]++] = (int32_t)e; }
            if (v != u && (v & adj_mask) == t) { nbr[fill[v]] = u; via[fill[v]++] = (int32_t)e; }
        }
    });
    std::vector<int64_t>().swap(fill);
    lap("adjacency");
    // ---- memo numbers of the edge dictionaries, relative to the end of the _node section: an edge's attribute dict
    // is memoised when the _adj section visits the edge for the first time, i.e. at its lower-numbered end
    std::vector<int64_t> first_before((size_t)n_nodes + 1, 0);         // first visits in the adjacency lists of earlier nodes
    parallel_for(n_nodes, n_threads, [&](int, int64_t i0, int64_t i1) {
        for (int64_t i = i0; i < i1; i++) {
            int64_t c = 0;
            for (int64_t a = start[i]; a < start[i + 1]; a++) c += nbr[a] >= i;
            first_before[i + 1] = c;
        }
    });
    for (int64_t i = 0; i < n_nodes; i++) first_before[i + 1] += first_before[i];
    std::vector<uint32_t> edge_rel((size_t)n_edges);
    parallel_for(n_nodes, n_threads, [&](int, int64_t i0, int64_t i1) {
        for (int64_t i = i0; i < i1; i++) {
            uint32_t m = (uint32_t)first_before[i];
            for (int64_t a = start[i]; a < start[i + 1]; a++) if (nbr[a] >= i) edge_rel[via[a]] = m++;
        }
    });
    std::vector<uint32_t> erel((size_t)start[n_nodes]);
    parallel_for(n_nodes, n_threads, [&](int, int64_t i0, int64_t i1) {
        for (int64_t i = i0; i < i1; i++)
            for (int64_t a = start[i]; a < start[i + 1]; a++) erel[a] = nbr[a] >= i ? RCQ_GRAPH_FIRST_VISIT : edge_rel[via[a]];
    });
    std::vector<int32_t>().swap(via); std::vector<uint32_t>().swap(edge_rel);
    lap("memo numbers");
    return rcq_graph_emit(path, n_samples, sample_names, n_nodes, node_sample.data(), node_gene.data(), start.data(), nbr.data(),
                          erel.data(), n_edges, counts);
}

extern "C" int rcq_graph_pickle(const char* path, int32_t n_samples, const char* const* sample_names,
                                int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_off,
                                const int32_t* sgene, const int32_t* qgene, int64_t n_isolated,
                                const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]) {
    try {   // no C++ exception may cross the C ABI
        std::vector<const int32_t*> sp((size_t)std::max(n_groups, 0), sgene), qp((size_t)std::max(n_groups, 0), qgene);   // one flat column: no bias
        return graph_pickle_impl(path, n_samples, sample_names, n_groups, g_s, g_q, g_off, (n_groups && sgene) ? sp.data() : nullptr,
                                 (n_groups && qgene) ? qp.data() : nullptr, n_isolated, iso_sample, iso_gene, counts);
    } catch (const std::bad_alloc&) {
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle: out of host memory");
    }
}

extern "C" int rcq_graph_pickle_groups(const char* path, int32_t n_samples, const char* const* sample_names,
                                       int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_len,
                                       const int32_t* const* sgene, const int32_t* const* qgene, int64_t n_isolated,
                                       const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]) {
    try {
        if (n_groups < 0 || (n_groups && (!g_len || !sgene || !qgene))) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_groups: bad argument");
        std::vector<int64_t> off((size_t)n_groups + 1, 0);
        std::vector<const int32_t*> sp((size_t)n_groups), qp((size_t)n_groups);
        for (int32_t g = 0; g < n_groups; g++) {
            if (g_len[g] < 0 || (g_len[g] && (!sgene[g] || !qgene[g]))) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle_groups: bad group");
            off[g + 1] = off[g] + g_len[g];
            sp[g] = sgene[g] - off[g]; qp[g] = qgene[g] - off[g];
        }
        return graph_pickle_impl(path, n_samples, sample_names, n_groups, g_s, g_q, off.data(), sp.data(), qp.data(), n_isolated,
                                 iso_sample, iso_gene, counts);
    } catch (const std::bad_alloc&) {
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle_groups: out of host memory");
    }
}
