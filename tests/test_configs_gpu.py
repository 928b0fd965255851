"""GPU: BASELINE.json configs[3] and configs[4] as parity cases (they are not bench lines).

configs[3] — simulated divergence sweep 0.01-0.20 (star tree with scaled root-to-tip lengths, reference
tree_gen.py:4-37), here with few genes per sample so the oracle finishes in seconds: every gene-matches
table, the restricted sums and the matrix equal the oracle's, and the distances grow with the
simulated divergence.
configs[4] — phase 2 only from 200 samples' precomputed tables (19,900 tables): bit-exact against the
numpy restatement of the reference's graph / ideal-component / Fraction code at a size it finishes in
seconds, then at 10^8 rows through properties of a construction whose answer is known in closed form."""
import itertools
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SWEEP = [0.005, 0.005, 0.02, 0.045, 0.075, 0.1, 0.1]   # root-to-tip; pairwise 0.01 .. 0.20 substitutions / site


def test_configs3_divergence_sweep_equals_the_oracle():
    from oracle import oracle as O
    from oracle import phase2_ref
    from rna_clique_b200 import _native
    from rna_clique_b200.engine import Engine, select_top_genes
    from rna_clique_b200.simulate import simulate, star_tree
    tree = star_tree(len(SWEEP), SWEEP)
    _, ts = simulate(len(SWEEP), 130, seed=31, mutation_rate=1.0, tree=tree)
    inputs, labels = [], []
    for t in ts:
        offs = t.offsets.numpy()
        inputs.append(dict(ascii=t.ascii.numpy(), starts=offs[:-1].copy(), lengths=np.diff(offs).astype(np.int32),
                           coverage=t.coverage, gene=t.gene, iso=t.iso))
        labels.append(f"od1/{t.name}_top.fasta")
    eng = Engine(0)
    res = eng.all_vs_all(inputs, labels, top_genes=100, keep_tables=True)
    osamples = []
    for s in inputs:
        keep = select_top_genes(s["coverage"], s["gene"], 100)
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
    exp = phase2_ref.phase2(len(SWEEP), otables)
    assert list(zip(res["nident_sum"].tolist(), res["denom_sum"].tolist())) == exp["sums"]
    _, m = phase2_ref.distance_matrix(labels, otables, exp["sums"])
    assert np.array_equal(m, res["matrix"])
    # the measured distance follows the simulated one: p = 3/4 (1 - exp(-4d/3)) under equal base frequencies,
    # slightly low at the far end where the X-drop trims the most diverged stretches
    order = {lab: k for k, lab in enumerate(res["labels"])}
    pairs = list(itertools.combinations(range(len(SWEEP)), 2))
    true_d = np.array([SWEEP[i] + SWEEP[j] for i, j in pairs])
    got_d = np.array([res["matrix"][order[labels[i]], order[labels[j]]] for i, j in pairs])
    p = 0.75 * (1 - np.exp(-4 * true_d / 3))
    assert np.abs(got_d - p).max() < 0.02 and got_d.min() > 0.005 and got_d.max() > 0.14
    far, near = got_d[true_d >= 0.15], got_d[true_d <= 0.05]
    assert far.min() > near.max()
    eng.close()


def _ortholog_tables(n_samples, n_genes, rng, noise=True):
    """Tables of ortholog groups: gene g of every sample matches gene g of every other, two rows per
    gene pair (forward and reverse hit), with drop-outs (a sample lacks the gene), missing edges and
    paralog rows (gene g of one sample also hits gene g' of another) when noise is set."""
    have = rng.random((n_samples, n_genes)) > (0.002 if noise else 0.0)
    tables = []
    for i in range(n_samples):
        for j in range(i + 1, n_samples):
            g = np.flatnonzero(have[i] & have[j]).astype(np.int32)
            if noise:
                g = g[rng.random(len(g)) > 2e-5]                      # a missing edge breaks completeness
            q = g.copy()
            if noise:
                wrong = rng.random(len(g)) < 2e-5
                q[wrong] = rng.integers(0, n_genes, size=int(wrong.sum()))   # paralog fan-out / merged components
            sg, qg = np.concatenate([g, g]), np.concatenate([q, q])
            length = rng.integers(300, 3000, size=len(sg)).astype(np.int32)
            gaps = rng.integers(0, 4, size=len(sg)).astype(np.int32)
            nident = (length - gaps - rng.integers(0, 50, size=len(sg))).astype(np.int32)
            tables.append(dict(s=i, q=j, sgene=sg, qgene=qg, nident=nident, length=length, gaps=gaps))
    return tables


def _run(ctx, n_samples, tables, stride):
    from rna_clique_b200 import _native
    row_off = np.concatenate([[0], np.cumsum([len(t["sgene"]) for t in tables])]).astype(np.int64)
    cat = lambda c: np.concatenate([t[c] for t in tables])
    return _native.phase2_host(ctx, n_samples, stride, row_off, [t["s"] for t in tables], [t["q"] for t in tables],
                               cat("sgene"), cat("qgene"), cat("nident"), cat("length"), cat("gaps"))


def test_configs4_phase2_of_200_samples_equals_the_oracle(ctx):
    from oracle import phase2_ref
    n_samples, n_genes = 200, 300
    tables = _ortholog_tables(n_samples, n_genes, np.random.default_rng(5))
    assert len(tables) == 19900
    exp = phase2_ref.phase2(n_samples, tables)
    res = _run(ctx, n_samples, tables, exp["stride"])
    c = res["counts"]
    assert (c["sample_count"], c["components"], c["ge_samples"], c["ideal"]) == (200, exp["n_components"], exp["n_ge_samples"], exp["n_ideal"])
    assert 0 < exp["n_ideal"] < n_genes and (c["nodes"], c["edges"]) == (exp["n_nodes"], exp["n_edges"])
    got = {(int(i // exp["stride"]), int(i % exp["stride"])) for i in np.flatnonzero(res["valid"])}
    assert got == exp["valid"]
    assert list(zip(res["nident_sum"].tolist(), res["denom_sum"].tolist())) == exp["sums"]
    names = [f"od1/S{k:03d}_top.fasta" for k in range(n_samples)]
    _, m = phase2_ref.distance_matrix(names, tables, exp["sums"])
    from rna_clique_b200.engine import Engine
    lab, mine = Engine.matrix_from_sums(names, [(t["s"], t["q"]) for t in tables], res["nident_sum"], res["denom_sum"])
    assert lab == sorted(names) and np.array_equal(mine, m)


def test_configs4_phase2_at_1e8_rows_has_the_closed_form_answer(ctx):
    """200 samples x 2,600 genes, no noise except that sample 7 lacks every gene divisible by 13: exactly the
    other genes are ideal, and each table's sums are plain masked column sums."""
    n_samples, n_genes = 200, 2600
    rng = np.random.default_rng(9)
    lacks = np.arange(n_genes) % 13 == 0
    tables = []
    for i in range(n_samples):
        for j in range(i + 1, n_samples):
            g = np.arange(n_genes, dtype=np.int32)
            if 7 in (i, j):
                g = g[~lacks]
            sg = np.concatenate([g, g])
            length = rng.integers(300, 3000, size=len(sg), dtype=np.int32)
            nident = length - rng.integers(0, 60, size=len(sg), dtype=np.int32)
            tables.append(dict(s=i, q=j, sgene=sg, qgene=sg, nident=nident, length=length, gaps=np.zeros(len(sg), dtype=np.int32)))
    rows = sum(len(t["sgene"]) for t in tables)
    assert rows > 1.0e8
    t0 = time.perf_counter()
    res = _run(ctx, n_samples, tables, n_genes)
    wall = time.perf_counter() - t0
    st = ctx.stage_times(reset=True)
    print(f"\nphase 2, {rows} rows in 19900 tables: {wall:.2f} s through rcq_phase2_host, stages {st}")
    c = res["counts"]
    assert c["ideal"] == int((~lacks).sum()) and c["components"] == n_genes and c["ge_samples"] == c["ideal"]
    valid = res["valid"].reshape(n_samples, n_genes)
    assert (valid == (~lacks)[None, :]).all()
    for k in (0, 6, 7 * 199 - 21, 12345, 19899):
        t = tables[k]
        keep = ~lacks[t["sgene"]]
        assert res["nident_sum"][k] == int(t["nident"][keep].astype(np.int64).sum())
        assert res["denom_sum"][k] == int(t["length"][keep].astype(np.int64).sum())
    assert res["nident_sum"].sum() == sum(int(t["nident"][~lacks[t["sgene"]]].astype(np.int64).sum()) for t in tables)
