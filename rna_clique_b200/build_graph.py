"""Gene-matches graph.  Mirrors the reference's build_graph.py:40-68: nodes are (sample path string,
gene id) tuples, one undirected edge per table row (repeated rows collapse).

The hot path never needs the networkx object — connected components and the ideal-component test
run on the device from flat (sample index, gene) arrays — so `EdgeArrays` keeps the flat form and
materialises / pickles as the reference's nx.Graph only on demand."""
from __future__ import annotations

from typing import Iterable

import networkx as nx
import numpy as np
import pandas as pd


class EdgeArrays:
    """Edges grouped per (ssample, qsample): lists of (ssample str, qsample str, sgene int32[], qgene int32[])."""

    def __init__(self):
        self.groups = []

    def add(self, ssample: str, qsample: str, sgene, qgene):
        self.groups.append((str(ssample), str(qsample), np.asarray(sgene, dtype=np.int32), np.asarray(qgene, dtype=np.int32)))

    def add_frame(self, df: pd.DataFrame):
        if len(df) == 0:
            return
        def only(col):   # the label, when the column holds one value (what find_homologs_and_save writes)
            c = df[col]
            if isinstance(c.dtype, pd.CategoricalDtype) and len(c.cat.categories) == 1 and not c.isna().any():
                return str(c.cat.categories[0])
            return None
        s1, q1 = only("ssample"), only("qsample")
        if s1 is not None and q1 is not None:
            self.add(s1, q1, df["sgene"].to_numpy(), df["qgene"].to_numpy())
            return
        ss, qs = df["ssample"].astype(str).to_numpy(), df["qsample"].astype(str).to_numpy()
        if (ss == ss[0]).all() and (qs == qs[0]).all():
            self.add(ss[0], qs[0], df["sgene"].to_numpy(), df["qgene"].to_numpy())
        else:  # a frame mixing sample pairs: split
            for (s, q), part in df.groupby([ss, qs], sort=False):
                self.add(s, q, part["sgene"].to_numpy(), part["qgene"].to_numpy())

    def to_networkx(self) -> nx.Graph:
        g = nx.Graph()
        for s, q, sg, qg in self.groups:
            g.add_edges_from(zip(zip([s] * len(sg), sg.tolist()), zip([q] * len(qg), qg.tolist())))
        return g

    def write_pickle(self, path) -> dict:
        """graph.pkl straight from the flat arrays (native writer, csrc/host_graph.cu): the same plain
        nx.Graph pickle as pickle.dump(self.to_networkx(), f), without building the Python object."""
        from . import _native
        index = {}
        for s, q, _, _ in self.groups:
            index.setdefault(s, len(index))
            index.setdefault(q, len(index))
        isolated = [(index.setdefault(str(n[0]), len(index)), int(n[1])) for n in getattr(self, "isolated", ())]
        return _native.graph_pickle(path, list(index), [(index[s], index[q], sg, qg) for s, q, sg, qg in self.groups], isolated)

    @classmethod
    def from_networkx(cls, g: nx.Graph) -> "EdgeArrays":
        buckets = {}
        for (s1, g1), (s2, g2) in g.edges():
            b = buckets.setdefault((s1, s2), ([], []))
            b[0].append(g1)
            b[1].append(g2)
        out = cls()
        for (s1, s2), (a, b) in buckets.items():
            out.add(s1, s2, a, b)
        out.isolated = [n for n in g.nodes if g.degree(n) == 0]
        return out


class GeneMatchGraph(nx.Graph):
    """nx.Graph that is filled lazily from EdgeArrays the first time networkx looks at it; it
    pickles as a plain networkx Graph (filtering_step.py:158-159 writes graph.pkl)."""

    def __init__(self, edges: EdgeArrays = None):
        super().__init__()
        self._rcq_edges = edges
        self._rcq_built = edges is None

    def _materialise(self):
        """Fills networkx's own dictionaries directly (same content and node order as add_edges_from:
        first appearance, u before v; one shared attribute dict per undirected edge).  Node tuples are
        canonical objects, so the pickle memoises each node once."""
        if not self._rcq_built:
            self._rcq_built = True
            node = object.__getattribute__(self, "_node")
            adj = object.__getattribute__(self, "_adj")
            canon = {}
            import gc
            gc_was_on = gc.isenabled()
            gc.disable()          # millions of small dicts: the cyclic collector only adds passes over them
            for s, q, sg, qg in self._rcq_edges.groups:
                cs, cq = canon.setdefault(s, {}), canon.setdefault(q, {})
                for a, b in zip(sg.tolist(), qg.tolist()):
                    u = cs.get(a)
                    if u is None:
                        u = cs[a] = (s, a)
                    v = cq.get(b)
                    if v is None:
                        v = cq[b] = (q, b)
                    du = adj.get(u)
                    if du is None:
                        du = adj[u] = {}
                        node[u] = {}
                    dv = adj.get(v)
                    if dv is None:
                        dv = adj[v] = {}
                        node[v] = {}
                    if v not in du:
                        du[v] = dv[u] = {}
            for n in getattr(self._rcq_edges, "isolated", ()):
                if n not in node:
                    adj[n] = {}
                    node[n] = {}
            if gc_was_on:
                gc.enable()
        return self

    @property
    def rcq_edges(self):
        """The flat edge arrays, or None once the graph has been modified through the networkx API."""
        return None if self.__dict__.get("_rcq_dirty") else self._rcq_edges

    def __reduce__(self):
        self._materialise()
        return (nx.Graph, (), {"graph": self.graph, "_node": self._node, "_adj": self._adj, "__networkx_cache__": {}})

    def __getattribute__(self, name):
        if name in ("_node", "_adj", "nodes", "edges", "adj", "degree") and not object.__getattribute__(self, "__dict__").get("_rcq_built", True):
            object.__getattribute__(self, "_materialise")()
        return object.__getattribute__(self, name)


def _mutator(name):
    def method(self, *args, **kwargs):
        self._materialise()
        self.__dict__["_rcq_dirty"] = True
        return getattr(nx.Graph, name)(self, *args, **kwargs)
    method.__name__ = name
    return method


for _name in ("add_node", "add_nodes_from", "remove_node", "remove_nodes_from", "add_edge", "add_edges_from",
              "add_weighted_edges_from", "remove_edge", "remove_edges_from", "update", "clear", "clear_edges"):
    setattr(GeneMatchGraph, _name, _mutator(_name))


def dump_graph(graph: nx.Graph, path) -> None:
    """pickle.dump(graph, f, HIGHEST_PROTOCOL) of filtering_step.py:158-159.  A GeneMatchGraph nobody has
    looked at through the networkx API yet is written natively from its edge arrays and stays lazy."""
    import pickle
    e = getattr(graph, "rcq_edges", None)
    if e is not None and not graph.__dict__.get("_rcq_built", True) and not graph.graph:
        e.write_pickle(path)
        return
    with open(path, "wb") as f:
        pickle.dump(graph, f, pickle.HIGHEST_PROTOCOL)


def build_graph(dfs: Iterable[pd.DataFrame]) -> nx.Graph:
    """build_graph(dfs) -> nx.Graph (build_graph.py:40-68).  Consumes the iterable (it is the
    unordered generator of find_all_pairs in the pipeline, filtering_step.py:157)."""
    e = EdgeArrays()
    for df in dfs:
        e.add_frame(df)
    return GeneMatchGraph(e)
