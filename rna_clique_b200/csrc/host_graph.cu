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
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>
#include "common.cuh"

namespace {
struct PickleOut {
    std::vector<uint8_t> b;
    uint32_t memo = 0;
    void op(uint8_t c) { b.push_back(c); }
    void raw(const void* p, size_t n) { const uint8_t* q = (const uint8_t*)p; b.insert(b.end(), q, q + n); }
    void u32(uint32_t v) { uint8_t x[4] = {(uint8_t)v, (uint8_t)(v >> 8), (uint8_t)(v >> 16), (uint8_t)(v >> 24)}; raw(x, 4); }
    uint32_t memoize() { op(0x94); return memo++; }                       // MEMOIZE
    void get(uint32_t id) { if (id < 256) { op('h'); op((uint8_t)id); } else { op('j'); u32(id); } }   // BINGET / LONG_BINGET
    void str(const char* s, size_t n) {                                    // SHORT_BINUNICODE / BINUNICODE (utf-8 bytes)
        if (n < 256) { op(0x8c); op((uint8_t)n); } else { op('X'); u32((uint32_t)n); }
        raw(s, n);
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
}  // namespace

// Groups g = 0..n_groups-1 hold rows [g_off[g], g_off[g+1]) of (sgene, qgene) between samples g_s[g] and
// g_q[g] (indices into sample_names); isolated nodes (degree 0) follow the edges.  counts[0] = nodes,
// counts[1] = distinct undirected edges, counts[2] = bytes written.
static int graph_pickle_impl(const char* path, int32_t n_samples, const char* const* sample_names,
                             int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_off,
                             const int32_t* sgene, const int32_t* qgene, int64_t n_isolated,
                             const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]) {
    if (!path || n_samples < 0 || n_groups < 0 || n_isolated < 0 || (n_samples && !sample_names) ||
        (n_groups && (!g_s || !g_q || !g_off)) || (n_isolated && (!iso_sample || !iso_gene)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: bad argument");
    const int64_t n_rows = n_groups ? g_off[n_groups] : 0;
    if (n_rows && (!sgene || !qgene)) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: bad argument");
    // ---- node ids in first-appearance order
    std::vector<int32_t> max_gene((size_t)n_samples, -1);
    for (int32_t g = 0; g < n_groups; g++) {
        if (g_s[g] < 0 || g_s[g] >= n_samples || g_q[g] < 0 || g_q[g] >= n_samples)
            return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: sample index out of range");
        for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) {
            if (sgene[r] < 0 || qgene[r] < 0) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: negative gene id");
            if (sgene[r] > max_gene[g_s[g]]) max_gene[g_s[g]] = sgene[r];
            if (qgene[r] > max_gene[g_q[g]]) max_gene[g_q[g]] = qgene[r];
        }
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
    // ---- distinct undirected edges in insertion order (open-addressing set of (min, max) node pairs)
    size_t cap = 1024;
    while (cap < 2 * (size_t)n_rows + 2) cap <<= 1;
    std::vector<uint64_t> set(cap, ~0ull);
    std::vector<int32_t> eu, ev;
    eu.reserve((size_t)n_rows); ev.reserve((size_t)n_rows);
    for (int32_t g = 0; g < n_groups; g++)
        for (int64_t r = g_off[g]; r < g_off[g + 1]; r++) {
            const int32_t u = node_of(g_s[g], sgene[r]), v = node_of(g_q[g], qgene[r]);
            const uint64_t key = u < v ? ((uint64_t)(uint32_t)u << 32) | (uint32_t)v : ((uint64_t)(uint32_t)v << 32) | (uint32_t)u;
            size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & (cap - 1);
            bool seen = false;
            while (set[h] != ~0ull) { if (set[h] == key) { seen = true; break; } h = (h + 1) & (cap - 1); }
            if (seen) continue;
            set[h] = key;
            eu.push_back(u); ev.push_back(v);
        }
    for (int64_t i = 0; i < n_isolated; i++) node_of(iso_sample[i], iso_gene[i]);
    const int64_t n_nodes = (int64_t)node_sample.size(), n_edges = (int64_t)eu.size();
    if (n_nodes + n_edges + n_samples + 16 >= (1ll << 32)) return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle: more than 2^32 memo entries");
    // ---- adjacency lists in edge-insertion order (CSR; a self-loop is listed once)
    std::vector<int64_t> start((size_t)n_nodes + 1, 0);
    for (int64_t e = 0; e < n_edges; e++) { start[eu[e] + 1]++; if (ev[e] != eu[e]) start[ev[e] + 1]++; }
    for (int64_t i = 0; i < n_nodes; i++) start[i + 1] += start[i];
    std::vector<int64_t> fill(start.begin(), start.end() - 1);
    std::vector<int32_t> nbr((size_t)start[n_nodes]);
    std::vector<int32_t> via((size_t)start[n_nodes]);
    for (int64_t e = 0; e < n_edges; e++) {
        nbr[fill[eu[e]]] = ev[e]; via[fill[eu[e]]++] = (int32_t)e;
        if (ev[e] != eu[e]) { nbr[fill[ev[e]]] = eu[e]; via[fill[ev[e]]++] = (int32_t)e; }
    }
    // ---- the stream
    PickleOut P;
    P.b.reserve((size_t)(64 + 16 * n_nodes + 10 * start[n_nodes]));
    P.op(0x80); P.op(4);                                                    // PROTO 4
    P.str("networkx.classes.graph", 22); P.str("Graph", 5); P.op(0x93);     // STACK_GLOBAL
    P.op(')'); P.op('R');                                                   // EMPTY_TUPLE REDUCE -> Graph()
    P.empty_dict(); P.mark();                                               // state = {
    P.str("graph", 5); P.empty_dict();
    std::vector<uint32_t> sample_memo((size_t)n_samples, 0xFFFFFFFFu), node_memo((size_t)n_nodes), edge_memo((size_t)n_edges, 0xFFFFFFFFu);
    P.str("_node", 5); P.empty_dict(); P.mark();
    for (int64_t i = 0; i < n_nodes; i++) {
        const int32_t s = node_sample[i];
        if (sample_memo[s] == 0xFFFFFFFFu) { P.str(sample_names[s], strlen(sample_names[s])); sample_memo[s] = P.memoize(); }
        else P.get(sample_memo[s]);
        P.integer(node_gene[i]);
        P.op(0x86);                                                         // TUPLE2
        node_memo[i] = P.memoize();
        P.empty_dict();
    }
    P.setitems();
    P.str("_adj", 4); P.empty_dict(); P.mark();
    for (int64_t i = 0; i < n_nodes; i++) {
        P.get(node_memo[i]);
        P.empty_dict();
        if (start[i + 1] > start[i]) {
            P.mark();
            for (int64_t a = start[i]; a < start[i + 1]; a++) {
                P.get(node_memo[nbr[a]]);
                uint32_t& m = edge_memo[via[a]];
                if (m == 0xFFFFFFFFu) { P.empty_dict(); m = P.memoize(); } else P.get(m);
            }
            P.setitems();
        }
    }
    P.setitems();
    P.str("__networkx_cache__", 18); P.empty_dict();
    P.setitems();                                                           // }
    P.op('b'); P.op('.');                                                   // BUILD STOP
    FILE* f = fopen(path, "wb");
    if (!f) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: cannot open the output file");
    const size_t w = fwrite(P.b.data(), 1, P.b.size(), f);
    if (fclose(f) != 0 || w != P.b.size()) return rcq_fail(RCQ_ERR_ARG, "rcq_graph_pickle: short write");
    if (counts) { counts[0] = n_nodes; counts[1] = n_edges; counts[2] = (int64_t)P.b.size(); }
    return 0;
}

extern "C" int rcq_graph_pickle(const char* path, int32_t n_samples, const char* const* sample_names,
                                int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_off,
                                const int32_t* sgene, const int32_t* qgene, int64_t n_isolated,
                                const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]) {
    try {   // no C++ exception may cross the C ABI
        return graph_pickle_impl(path, n_samples, sample_names, n_groups, g_s, g_q, g_off, sgene, qgene, n_isolated,
                                 iso_sample, iso_gene, counts);
    } catch (const std::bad_alloc&) {
        return rcq_fail(RCQ_ERR_CAPACITY, "rcq_graph_pickle: out of host memory");
    }
}
