"""GPU: parity at BASELINE.json configs[0] size against the oracle, and size-independent properties
at configs[1] transcript counts (order independence, stream-count independence, matrix shape)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(taxa, count, seed, **kw):
    from rna_clique_b200.simulate import simulate
    _, ts = simulate(taxa, count, seed=seed, **kw)
    inputs, labels = [], []
    for t in ts:
        offs = t.offsets.numpy()
        inputs.append(dict(ascii=t.ascii.numpy(), starts=offs[:-1].copy(), lengths=np.diff(offs).astype(np.int32),
                           coverage=t.coverage, gene=t.gene, iso=t.iso))
        labels.append(f"od1/{t.name}_top.fasta")
    return inputs, labels


def test_configs0_tables_and_matrix_equal_the_oracle():
    """3 simulated samples x 5,000 transcripts, top_genes = 2,000 (the reference's PR1 run):
    every gene-matches table, the restricted sums and the fp64 matrix, bit for bit."""
    from oracle import oracle as O
    from oracle import phase2_ref
    from rna_clique_b200 import _native
    from rna_clique_b200.engine import Engine, select_top_genes
    inputs, labels = _inputs(3, 5000, 487)
    eng = Engine(0)
    res = eng.all_vs_all(inputs, labels, top_genes=2000, keep_tables=True)
    osamples = []
    for s in inputs:
        keep = select_top_genes(s["coverage"], s["gene"], 2000)
        assert len(keep) == 2000
        offs = np.concatenate([[0], np.cumsum(s["lengths"][keep])]).astype(np.int64)
        buf = np.concatenate([s["ascii"][a:a + n] for a, n in zip(s["starts"][keep], s["lengths"][keep])])
        osamples.append(O.Sample(buf, offs, s["gene"][keep], s["iso"][keep]))
    otables = []
    for (i, j), t in zip(res["pairs"], res["tables"]):
        want = O.search_pair(osamples[i], osamples[j]).rows(2)
        cols = t.download()
        got = np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)
        assert np.array_equal(got, want), (i, j)
        otables.append(dict(s=i, q=j, sgene=want[:, 12], qgene=want[:, 10], nident=want[:, 9], length=want[:, 0], gaps=want[:, 8]))
    exp = phase2_ref.phase2(3, otables)
    assert list(zip(res["nident_sum"].tolist(), res["denom_sum"].tolist())) == exp["sums"]
    _, m = phase2_ref.distance_matrix(labels, otables, exp["sums"])
    assert np.array_equal(m, res["matrix"])
    assert res["counts"]["ideal"] == exp["n_ideal"] > 1900
    eng.close()


def test_full_width_properties_do_not_depend_on_order_or_streams():
    """4 samples x 60,000 transcripts, top_genes = 50,000 (configs[1] per-sample size): the matrix is
    symmetric with a zero diagonal, identical for 1 and 3 streams, and (to 1e-6) independent of the
    input order; every pair contributes exactly 2 x 50,000 rows and every gene is an ideal component."""
    from rna_clique_b200.engine import Engine
    inputs, labels = _inputs(4, 60000, 487)
    e3, e1 = Engine(0, streams=3), Engine(0, streams=1)
    a = e3.all_vs_all(inputs, labels, top_genes=50000)
    b = e1.all_vs_all(inputs, labels, top_genes=50000)
    perm = [2, 0, 3, 1]
    c = e3.all_vs_all([inputs[k] for k in perm], [labels[k] for k in perm], top_genes=50000)
    m = a["matrix"]
    assert m.shape == (4, 4) and np.array_equal(m, m.T) and (np.diag(m) == 0).all() and (m[~np.eye(4, dtype=bool)] > 0).all()
    assert np.array_equal(m, b["matrix"]) and a["stats"] == b["stats"]
    # rows/cols are sorted by label.  Swapping which sample of a pair plays transcripts1 re-runs the greedy
    # extension with query and subject exchanged; its tie rules (mismatch before query gap before subject
    # gap) are not symmetric, so a handful of alignments trade two mismatches for a gap pair: the
    # matrices agree to ~2e-8 (measured), not bit for bit (the reference's blastn runs have the same property).
    assert c["labels"] == a["labels"] and np.abs(c["matrix"] - m).max() < 1e-6
    assert a["stats"]["n_rows"] == 6 * 2 * 50000 and a["counts"]["ideal"] == 50000 == a["counts"]["components"]
    assert a["stats"]["n_fwd"] == a["stats"]["n_bwd"] == 6 * 50000
    e3.close(); e1.close()


def test_subset_rerun_on_cached_tables_equals_a_fresh_run():
    """make_subset semantics: phase 2 restricted to a subset of the samples, from the device tables of the
    full run, equals running the subset from scratch (and the numpy oracle on the downloaded tables)."""
    from oracle import phase2_ref
    from rna_clique_b200.engine import Engine
    inputs, labels = _inputs(5, 1500, 12, mutation_rate=0.03, indel_rate=0.002)
    eng = Engine(0)
    full = eng.all_vs_all(inputs, labels, top_genes=1000, keep_tables=True)
    keep = [0, 2, 3]
    sub = eng.subset_distances(full, labels, keep)
    fresh = eng.all_vs_all([inputs[k] for k in keep], [labels[k] for k in keep], top_genes=1000)
    assert sub["labels"] == fresh["labels"] and np.array_equal(sub["matrix"], fresh["matrix"])
    assert sub["counts"]["sample_count"] == 3 and sub["counts"]["ideal"] == fresh["counts"]["ideal"]
    otables = []
    for (i, j), t in zip(full["pairs"], full["tables"]):
        if i in keep and j in keep:
            c = t.download()
            otables.append(dict(s=i, q=j, sgene=c["sgene"], qgene=c["qgene"], nident=c["nident"], length=c["length"], gaps=c["gaps"]))
    exp = phase2_ref.phase2(5, otables)
    assert list(zip(sub["nident_sum"].tolist(), sub["denom_sum"].tolist())) == exp["sums"]
    # the full run has fewer ideal components than the subset never had more samples to satisfy
    assert full["counts"]["ideal"] <= sub["counts"]["ideal"]
    eng.close()
