"""graph.pkl written natively from the edge arrays (csrc/host_graph.cu) against the object networkx builds
from the same rows (build_graph.py:40-68 + filtering_step.py:158-159): same class, same nodes and
neighbours in the same order, repeated rows collapsed, one shared attribute dict per undirected edge."""
import pickle

import networkx as nx
import numpy as np
import pandas as pd

from rna_clique_b200.build_graph import EdgeArrays, GeneMatchGraph, build_graph, dump_graph


def _edges(rng, n_samples=5, n_genes=400, big=False):
    names = [f"/data/out/od1/sample_{k}_top.fasta" for k in range(n_samples)]
    if big:
        names[1] = "/data/" + "x" * 300 + "/s1_top.fasta"      # a path longer than 255 bytes (BINUNICODE)
    e = EdgeArrays()
    for i in range(n_samples):
        for j in range(i + 1, n_samples):
            g = rng.permutation(n_genes)[: n_genes * 3 // 4].astype(np.int32)
            q = g.copy()
            wrong = rng.random(len(g)) < 0.05
            q[wrong] = rng.integers(0, n_genes, size=int(wrong.sum()))
            dup = rng.random(len(g)) < 0.3                        # tie rows repeat a gene pair
            e.add(names[i], names[j], np.concatenate([g, g[dup]]), np.concatenate([q, q[dup]]))
    return e


def _reference_graph(e):
    g = nx.Graph()
    for s, q, sg, qg in e.groups:
        g.add_edges_from(zip(zip([s] * len(sg), sg.tolist()), zip([q] * len(qg), qg.tolist())))
    g.add_nodes_from(getattr(e, "isolated", ()))
    return g


def _same(a, b):
    assert type(a) is nx.Graph and a.graph == {} == b.graph
    assert list(a.nodes) == list(b.nodes)
    assert all(list(a.adj[n]) == list(b.adj[n]) for n in b.nodes)
    assert all(a.nodes[n] == {} for n in a.nodes)
    assert a.number_of_edges() == b.number_of_edges()
    for u, v in list(a.edges)[:200]:
        assert a._adj[u][v] is a._adj[v][u] and a._adj[u][v] == {}


def test_native_pickle_equals_networkx(tmp_path):
    rng = np.random.default_rng(3)
    for big in (False, True):
        e = _edges(rng, big=big)
        e.isolated = [(e.groups[0][0], 100000), ("/somewhere/else_top.fasta", 7)]   # a gene id needing 4 bytes
        info = e.write_pickle(tmp_path / "graph.pkl")
        with open(tmp_path / "graph.pkl", "rb") as f:
            got = pickle.load(f)
        want = _reference_graph(e)
        _same(got, want)
        assert info["nodes"] == want.number_of_nodes() and info["edges"] == want.number_of_edges()
        # usable like any other graph afterwards
        got.add_edge(("a", 1), ("b", 2))
        assert nx.number_connected_components(got) == nx.number_connected_components(want) + 1


def test_dump_graph_keeps_the_graph_lazy_and_falls_back(tmp_path):
    rng = np.random.default_rng(4)
    e = _edges(rng, n_samples=3, n_genes=50)
    frames = []
    for s, q, sg, qg in e.groups:
        frames.append(pd.DataFrame({"ssample": s, "qsample": q, "sgene": sg, "qgene": qg}))
    g = build_graph(iter(frames))
    dump_graph(g, tmp_path / "a.pkl")
    assert isinstance(g, GeneMatchGraph) and not g.__dict__["_rcq_built"]          # never materialised
    a = pickle.load(open(tmp_path / "a.pkl", "rb"))
    _same(a, _reference_graph(e))
    g.add_node(("extra", 1))                                                        # now a modified networkx object
    dump_graph(g, tmp_path / "b.pkl")
    b = pickle.load(open(tmp_path / "b.pkl", "rb"))
    assert ("extra", 1) in b and b.number_of_edges() == a.number_of_edges()
    # empty graph and self-loop
    empty = EdgeArrays()
    empty.write_pickle(tmp_path / "e.pkl")
    assert pickle.load(open(tmp_path / "e.pkl", "rb")).number_of_nodes() == 0
    loop = EdgeArrays()
    loop.add("s", "s", [5, 5], [5, 6])
    loop.write_pickle(tmp_path / "l.pkl")
    lg = pickle.load(open(tmp_path / "l.pkl", "rb"))
    assert list(lg.edges) == [(("s", 5), ("s", 5)), (("s", 5), ("s", 6))] and list(lg.adj[("s", 5)]) == [("s", 5), ("s", 6)]
    # large sparse gene ids (a custom id parser): the hash-map node table
    sp = EdgeArrays()
    sp.add("a", "b", [2_000_000_000, 5, 2_000_000_000], [1_999_999_999, 5, 1_999_999_999])
    sp.write_pickle(tmp_path / "s.pkl")
    sg = pickle.load(open(tmp_path / "s.pkl", "rb"))
    assert list(sg.nodes) == [("a", 2_000_000_000), ("b", 1_999_999_999), ("a", 5), ("b", 5)] and sg.number_of_edges() == 2


def test_sorted_tables_parallel_parts_and_repeated_sample_pairs(tmp_path):
    """The writer's three duplicate-row paths and its parallel adjacency section: tables sorted by gene pair with tie
    rows next to each other (one comparison per row), an unsorted table (a hash set of its own), two groups
    joining the same samples (one global set), on enough nodes for several parts."""
    rng = np.random.default_rng(9)
    names = [f"od1/S{k}_top.fasta" for k in range(4)]
    n_genes = 3000
    e = EdgeArrays()
    for i in range(4):
        for j in range(i + 1, 4):
            q = np.sort(rng.choice(n_genes, size=n_genes * 4 // 5, replace=False)).astype(np.int32)
            s = q.copy()
            wrong = rng.random(len(q)) < 0.03
            s[wrong] = rng.integers(0, n_genes, size=int(wrong.sum()))
            rep = np.repeat(np.arange(len(q)), 1 + (rng.random(len(q)) < 0.5))     # tie rows directly after their twin
            e.add(names[i], names[j], s[rep], q[rep])
    got_info = e.write_pickle(tmp_path / "sorted.pkl")
    _same(pickle.load(open(tmp_path / "sorted.pkl", "rb")), _reference_graph(e))
    assert got_info["nodes"] > 4096
    # one table out of order
    s0, q0, sg, qg = e.groups[2]
    perm = rng.permutation(len(sg))
    e.groups[2] = (s0, q0, sg[perm], qg[perm])
    e.write_pickle(tmp_path / "mixed.pkl")
    _same(pickle.load(open(tmp_path / "mixed.pkl", "rb")), _reference_graph(e))
    # the same sample pair twice, in both orientations, with edges in common
    s0, q0, sg, qg = e.groups[0]
    e.add(q0, s0, qg[:500], sg[:500])
    e.add(s0, q0, sg[200:900], qg[200:900])
    e.write_pickle(tmp_path / "repeat.pkl")
    _same(pickle.load(open(tmp_path / "repeat.pkl", "rb")), _reference_graph(e))
