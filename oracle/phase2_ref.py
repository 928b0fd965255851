"""CPU ORACLE (test infrastructure, NOT the product) for phase 2: gene-match graph, ideal
components, restricted per-pair sums and the distance matrix.

numpy/scipy restatement of the reference's
  build_graph.py:40-68           nodes (sample, gene); one undirected edge per table row
  filtered_distance.py:25-39     ideal component <=> |V| == #samples and 2|E| == |V|(|V|-1)
  filtered_distance.py:171-182   #samples = distinct samples among graph nodes (unless given)
  filtered_distance.py:184-210   valid genes; a row survives iff both of its genes are valid
  similarity_computer.py:21-42   Fraction(sum nident, sum length - sum gaps) per table
  similarity_computer.py:216-251,285-312  diagonal 1, d = 1 - sim (exact), float() last, samples sorted
Pinned against the reference's own code by tests/test_phase2_oracle.py on tests/golden/phase2_*.json."""
from fractions import Fraction

import numpy as np
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import connected_components


def phase2(n_samples, tables, sample_count=None):
    """tables: list of dict(s=int sample index, q=int, sgene, qgene, nident, length, gaps arrays).
    Returns dict(valid=set of (sample, gene), sums=[(nident_sum, denom_sum)], sample_count, n_ideal)."""
    stride = 1 + max([0] + [int(max(t["sgene"].max(initial=0), t["qgene"].max(initial=0))) for t in tables])
    us, vs = [], []
    for t in tables:
        us.append(t["s"] * stride + np.asarray(t["sgene"], dtype=np.int64))
        vs.append(t["q"] * stride + np.asarray(t["qgene"], dtype=np.int64))
    u = np.concatenate(us) if us else np.zeros(0, dtype=np.int64)
    v = np.concatenate(vs) if vs else np.zeros(0, dtype=np.int64)
    nodes, inv = np.unique(np.concatenate([u, v]), return_inverse=True)
    ui, vi = inv[:len(u)], inv[len(u):]
    n = len(nodes)
    if sample_count is None:
        sample_count = len(np.unique(nodes // stride))
    if n:
        _, comp = connected_components(coo_matrix((np.ones(len(ui)), (ui, vi)), shape=(n, n)), directed=False)
    else:
        comp = np.zeros(0, dtype=np.int64)
    vcount = np.bincount(comp, minlength=comp.max(initial=-1) + 1)
    lo, hi = np.minimum(ui, vi), np.maximum(ui, vi)
    uniq = np.unique(lo * max(n, 1) + hi)
    ecount = np.bincount(comp[uniq // max(n, 1)], minlength=len(vcount))
    ideal = (vcount == sample_count) & (2 * ecount == vcount * (vcount - 1))
    valid_nodes = nodes[ideal[comp]] if n else nodes
    valid = set(zip((valid_nodes // stride).tolist(), (valid_nodes % stride).tolist()))
    vset = np.zeros(n_samples * stride, dtype=bool)
    vset[valid_nodes] = True
    sums = []
    for t, uu, vv in zip(tables, us, vs):
        keep = vset[uu] & vset[vv]
        ni = int(np.asarray(t["nident"], dtype=np.int64)[keep].sum())
        de = int(np.asarray(t["length"], dtype=np.int64)[keep].sum() - np.asarray(t["gaps"], dtype=np.int64)[keep].sum())
        sums.append((ni, de))
    return dict(valid=valid, sums=sums, sample_count=int(sample_count), n_ideal=int(ideal.sum()), stride=stride,
                n_nodes=n, n_edges=len(uniq), n_components=int(len(vcount)),
                n_ge_samples=int((vcount >= sample_count).sum()))   # plot_component_sizes.py:214-230 statistics


def distance_matrix(names, tables, sums):
    """names: sample label per index; returns (sorted labels, fp64 matrix) exactly as the reference:
    Fraction until the final float()."""
    present = sorted({names[t["s"]] for t in tables} | {names[t["q"]] for t in tables})
    pos = {s: i for i, s in enumerate(present)}
    m = np.zeros((len(present), len(present)))
    for t, (ni, de) in zip(tables, sums):
        if de == 0:
            raise ZeroDivisionError("a restricted table is empty")
        d = float(1 - Fraction(ni, de))
        a, b = pos[names[t["s"]]], pos[names[t["q"]]]
        m[a, b] = m[b, a] = d
    return present, m
