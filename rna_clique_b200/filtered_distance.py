"""SampleSimilarity: ideal-component filter + distance matrix.  Mirrors the reference's
filtered_distance.py:129-317 and the base class similarity_computer.py:44-421, with the networkx
component walk (filtered_distance.py:25-39), the two pandas merges per table (:66-124) and the
column sums (similarity_computer.py:21-42) replaced by ONE call into the CUDA library
(rcq_phase2_host: union-find components, ideal test, validity bitmap, masked int64 sums).
Rationals stay exact until the final float() (similarity_computer.py:305)."""
from __future__ import annotations

import pickle
from collections.abc import Mapping
from fractions import Fraction
from functools import cached_property
from pathlib import Path
from typing import Iterable, Iterator, Optional

import networkx as nx
import numpy as np
import pandas as pd

from . import _native
from .build_graph import EdgeArrays, GeneMatchGraph
from .gene_matches_tables import read_table


class NoIdealComponentsError(Exception):
    """Raised when a gene-matches table has no row inside ideal components
    (reference filtered_distance.py:126-127,242-247)."""


class PairDict(dict):
    """Unordered-pair mapping: d[[a, b]], d[(a, b)] and d[frozenset((a, b))] all work, and (a, a)
    is a valid key (the role multiset_key_dict.MultisetKeyDict plays upstream)."""

    @staticmethod
    def _k(key):
        k = sorted(key)
        return tuple(k * 2 if len(k) == 1 else k)

    def __getitem__(self, key):
        return super().__getitem__(self._k(key))

    def __setitem__(self, key, value):
        super().__setitem__(self._k(key), value)

    def __contains__(self, key):
        return super().__contains__(self._k(key))

    def key_elements(self):
        return {e for k in self.keys() for e in k}


_default_ctx: Optional[_native.Context] = None


def _ctx() -> _native.Context:
    global _default_ctx
    if _default_ctx is None:
        from .find_homologs import default_cache
        _default_ctx = default_cache().ctx
    return _default_ctx


class SampleSimilarity:
    sample_gene_columns = [["ssample", "sgene"], ["qsample", "qgene"]]

    def __init__(self, graph: Optional[nx.Graph], comparison_dfs: Iterable, sample_count: Optional[int] = None,
                 ctx: Optional[_native.Context] = None, _device: Optional[dict] = None):
        """_device (set by rna_clique()): the device-resident tables the graph was built from
        (filtering_step); while the graph is untouched the reduction runs on them (rcq_phase2_tables) and
        `comparison_dfs` - lazy readers of the same tables on disk - is not consumed."""
        self.graph = graph
        self._sample_count = sample_count
        self._source = comparison_dfs
        self._stored = comparison_dfs if isinstance(comparison_dfs, Mapping) else None
        self._ctx_override = ctx
        self._device = _device

    # ---- ingestion --------------------------------------------------------------------------
    @staticmethod
    def mapping_from_dfs(dfs: Iterable[pd.DataFrame]) -> Iterator[tuple[frozenset, pd.DataFrame]]:
        """(frozenset{qsample, ssample}, df) per table (similarity_computer.py:90-115)."""
        for df in dfs:
            yield frozenset((df["qsample"][0], df["ssample"][0])), df

    @staticmethod
    def _read_table(path: Path) -> pd.DataFrame:
        df = read_table(Path(path))
        return df.drop(columns=[c for c in ("qseqid", "sseqid") if c in df.columns])

    @property
    def comparison_dfs(self):
        return self._stored

    @cached_property
    def _tables(self):
        """Compact per-table columns, consumed once from the (possibly lazy) source."""
        it = self._source.items() if isinstance(self._source, Mapping) else self._source
        out = []
        for _, df in it:
            if len(df) == 0:
                raise KeyError(0)   # the reference reads df["qsample"][0] of every table
            out.append(dict(ssample=str(df["ssample"].iloc[0]), qsample=str(df["qsample"].iloc[0]),
                            **{c: df[c].to_numpy().astype(np.int32) for c in ("sgene", "qgene", "nident", "length", "gaps")}))
        return out

    # ---- the device call ----------------------------------------------------------------------
    def _use_device(self) -> bool:
        return (self._device is not None and isinstance(self.graph, GeneMatchGraph) and self.graph.rcq_edges is not None
                and all(t.h is not None for t in self._device["tables"]))

    @cached_property
    def _pair_names(self):
        """(ssample, qsample) of every comparison table, in table order."""
        if self._use_device():
            d = self._device
            if any(d["empty"]):
                raise KeyError(0)   # the reference reads df["qsample"][0] of every table
            return [(d["labels"][i], d["labels"][j]) for i, j in d["pairs"]]
        return [(t["ssample"], t["qsample"]) for t in self._tables]

    def _phase2_device(self):
        d = self._device
        names = sorted({s for pair in self._pair_names for s in pair})
        idx = {s: i for i, s in enumerate(names)}
        eng = d["engine"]
        stride = 1 + max([0] + [int(max(sg.max(initial=0), qg.max(initial=0))) for _, _, sg, qg in self.graph.rcq_edges.groups])
        res = _native.phase2_tables(self._ctx_override or eng.ctx, len(names), stride, d["tables"],
                                    [idx[d["labels"][i]] for i, _ in d["pairs"]], [idx[d["labels"][j]] for _, j in d["pairs"]],
                                    self._sample_count, True)
        res.update(names=names, stride=stride, index=idx, n_real=len(d["tables"]))
        return res

    @cached_property
    def _phase2(self):
        if self._use_device():
            return self._phase2_device()
        tabs = self._tables
        edges = None
        extra_samples = set()
        if isinstance(self.graph, GeneMatchGraph) and self.graph.rcq_edges is not None:
            edges = self.graph.rcq_edges       # flat arrays, still exact: the graph was not modified
        elif self.graph is not None:
            edges = EdgeArrays.from_networkx(self.graph)
            extra_samples = {n[0] for n in getattr(edges, "isolated", [])}
        names = sorted({t["ssample"] for t in tabs} | {t["qsample"] for t in tabs} | extra_samples |
                       ({g[0] for g in edges.groups} | {g[1] for g in edges.groups} if edges else set()))
        idx = {s: i for i, s in enumerate(names)}
        rows = [(idx[t["ssample"]], idx[t["qsample"]], t["sgene"], t["qgene"], t["nident"], t["length"], t["gaps"],
                 2 if edges is not None else 3) for t in tabs]
        if edges is not None:
            for s, q, sg, qg in edges.groups:
                z = np.zeros(len(sg), dtype=np.int32)
                rows.append((idx[s], idx[q], sg, qg, z, z, z, 1))
        stride = 1 + max([0] + [int(max(r[2].max(initial=0), r[3].max(initial=0))) for r in rows])
        cat = lambda k: np.concatenate([r[k] for r in rows]) if rows else np.zeros(0, dtype=np.int32)
        row_off = np.concatenate([[0], np.cumsum([len(r[2]) for r in rows])]).astype(np.int64)
        sc = self._sample_count
        if sc is None and extra_samples:
            # samples are counted over the graph's nodes only (filtered_distance.py:171-182): its edges and its
            # isolated nodes; a sample that appears in the tables alone does not count
            sc = len(extra_samples | {g[0] for g in edges.groups} | {g[1] for g in edges.groups})
        res = _native.phase2_host(self._ctx_override or _ctx(), len(names), stride, row_off, [r[0] for r in rows],
                                  [r[1] for r in rows], cat(2), cat(3), cat(4), cat(5), cat(6), sc, True,
                                  t_flags=[r[7] for r in rows])
        res.update(names=names, stride=stride, index=idx, n_real=len(tabs))
        return res

    # ---- reference API ------------------------------------------------------------------------
    @property
    def sample_count(self) -> int:
        if self._sample_count is None:
            self._sample_count = int(self._phase2["counts"]["sample_count"])
        return self._sample_count

    @cached_property
    def valid(self) -> pd.DataFrame:
        p = self._phase2
        v = np.flatnonzero(p["valid"])
        return pd.DataFrame({"sample": np.asarray(p["names"], dtype=object)[v // p["stride"]] if len(v) else [],
                             "gene": (v % p["stride"]).astype(np.int64)}, columns=["sample", "gene"])

    @property
    def ideal_component_count(self) -> int:
        return int(self._phase2["counts"]["ideal"])

    def component_statistics(self):
        """(samples, components, components with >= samples nodes, ideal components): the four numbers
        `plot_component_sizes --statistics` prints (plot_component_sizes.py:214-230), from the union-find
        counts of the device pass."""
        c = self._phase2["counts"]
        return int(c["sample_count"]), int(c["components"]), int(c["ge_samples"]), int(c["ideal"])

    def restricted(self, comp_df: pd.DataFrame) -> pd.DataFrame:
        p = self._phase2
        if len(comp_df) == 0:
            return comp_df
        keep = np.ones(len(comp_df), dtype=bool)
        for scol, gcol in self.sample_gene_columns:
            si = comp_df[scol].astype(str).map(p["index"]).to_numpy()
            g = comp_df[gcol].to_numpy().astype(np.int64)
            ok = ~pd.isna(si) & (g < p["stride"]) & (g >= 0)
            node = np.where(ok, np.nan_to_num(si.astype(float)).astype(np.int64) * p["stride"] + np.where(ok, g, 0), 0)
            keep &= ok & (p["valid"][node] != 0)
        return comp_df[keep]

    def restricted_comparison_dfs(self):
        if self._stored is None:
            raise ValueError("comparison tables were not stored (store_dfs=False)")
        for k, df in self._stored.items():
            yield k, self.restricted(df)

    @cached_property
    def similarities(self) -> PairDict:
        p = self._phase2
        d = PairDict()
        for (ss, qs), ni, de in zip(self._pair_names, p["nident_sum"].tolist(), p["denom_sum"].tolist()):
            if de == 0:
                raise NoIdealComponentsError()
            d[(qs, ss)] = Fraction(int(ni), int(de))
        self._samples = sorted(d.key_elements())
        for s in self._samples:
            d[(s, s)] = 1
        return d

    @property
    def samples(self):
        self.similarities  # noqa: B018
        return self._samples

    @staticmethod
    def similarity_to_dissimilarity(sim):
        return 1 - sim

    def get_similarities(self) -> PairDict:
        return self.similarities

    def get_dissimilarities(self) -> PairDict:
        d = PairDict()
        for k, v in self.similarities.items():
            d[k] = 1 - v
        return d

    def _pair_dict_to_matrix(self, d: PairDict) -> np.ndarray:
        rows = [[float(d[[a, b]]) for b in self.samples] for a in self.samples]
        return np.vstack(rows) if rows else np.zeros(shape=(0, 0))

    def _matrix_to_df(self, mat: np.ndarray) -> pd.DataFrame:
        return pd.DataFrame(mat).set_axis(self.samples, axis=1).set_axis(self.samples, axis=0)

    def get_similarity_matrix(self) -> np.ndarray:
        return self._pair_dict_to_matrix(self.get_similarities())

    def get_dissimilarity_matrix(self) -> np.ndarray:
        return self._pair_dict_to_matrix(self.get_dissimilarities())

    def get_similarity_df(self) -> pd.DataFrame:
        return self._matrix_to_df(self.get_similarity_matrix())

    def get_dissimilarity_df(self) -> pd.DataFrame:
        return self._matrix_to_df(self.get_dissimilarity_matrix())

    @classmethod
    def from_filenames(cls, graph_fn: Path, comparison_fns: Iterable[Path], store_dfs: bool = True,
                       sample_count: Optional[int] = None, **kwargs) -> "SampleSimilarity":
        """Phase-2-only entry (filtered_distance.py:249-317): graph pickle + table files."""
        with open(graph_fn, "rb") as f:
            graph = pickle.load(f)
        pairs = cls.mapping_from_dfs(cls._read_table(p) for p in comparison_fns)
        if store_dfs:
            pairs = dict(pairs)
        return cls(graph, pairs, sample_count, **kwargs)
