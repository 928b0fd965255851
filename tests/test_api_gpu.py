"""GPU: the reference-facing Python API (rna_clique(), SampleSimilarity, HomologFinder, find_all_pairs)
end to end through the C ABI, against the oracle and the reference-generated fixtures."""
import pickle

import networkx as nx
import numpy as np
import pandas as pd
import pytest

from test_phase2_oracle import CASES, load_case

pytestmark = pytest.mark.gpu


def _frames(names, tables):
    out = []
    for t in tables:
        df = pd.DataFrame({c: t[c] for c in ("sgene", "qgene", "nident", "length", "gaps")})
        df["ssample"], df["qsample"] = names[t["s"]], names[t["q"]]
        out.append(df)
    return out


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("graph_kind", ["lazy", "networkx"])
def test_sample_similarity_against_reference_fixtures(ctx, name, graph_kind):
    from rna_clique_b200.build_graph import build_graph
    from rna_clique_b200.filtered_distance import NoIdealComponentsError, SampleSimilarity
    rec, names, tables = load_case(name)
    exp = rec["expect"]
    dfs = _frames(names, tables)
    graph = build_graph(dfs)
    if graph_kind == "networkx":
        graph = pickle.loads(pickle.dumps(graph))
        assert type(graph) is nx.Graph
    sim = SampleSimilarity(graph, SampleSimilarity.mapping_from_dfs(iter(dfs)), rec.get("sample_count"), ctx=ctx)
    assert sim.sample_count == exp["sample_count"]
    assert sorted([s, int(g)] for s, g in sim.valid.itertuples(index=False)) == exp["valid"]
    if exp["error"]:
        with pytest.raises(NoIdealComponentsError):
            sim.get_dissimilarity_df()
        return
    df = sim.get_dissimilarity_df()
    assert list(df.index) == exp["samples"] and list(df.columns) == exp["samples"]
    assert np.array_equal(df.to_numpy(), np.asarray(exp["matrix"]))       # fp64, bit for bit
    # restricted(): same rows the two validity probes keep
    v = {(s, int(g)) for s, g in sim.valid.itertuples(index=False)}
    for d in dfs[:3]:
        keep = [(s, g) in v and (q, h) in v for s, g, q, h in zip(d["ssample"], d["sgene"], d["qsample"], d["qgene"])]
        assert sim.restricted(d).equals(d[np.asarray(keep, dtype=bool)])


def test_graph_that_is_not_the_union_of_the_tables(ctx):
    """SampleSimilarity takes graph and tables separately (filtered_distance.py:162-169): the valid set
    comes from the graph alone."""
    from rna_clique_b200.filtered_distance import SampleSimilarity
    mk = lambda s, q, sg, qg: pd.DataFrame(dict(ssample=s, qsample=q, sgene=sg, qgene=qg, nident=[90] * len(sg),
                                                 length=[100] * len(sg), gaps=[0] * len(sg)))
    dfs = [mk("a", "b", [1, 2], [1, 2]), mk("a", "c", [1, 2], [1, 2]), mk("b", "c", [1, 2], [1, 2])]
    g = nx.Graph()
    for d in dfs:
        for s, q, sg, qg in zip(d["ssample"], d["qsample"], d["sgene"], d["qgene"]):
            if sg == 1:                      # the graph only knows gene 1
                g.add_edge((s, int(sg)), (q, int(qg)))
    sim = SampleSimilarity(g, SampleSimilarity.mapping_from_dfs(iter(dfs)), ctx=ctx)
    assert sorted(sim.valid.itertuples(index=False, name=None)) == [("a", 1), ("b", 1), ("c", 1)]
    assert sim.get_dissimilarity_df().loc["a", "b"] == 0.1      # float(1 - Fraction(90, 100))


def test_rna_clique_end_to_end(tmp_path):
    from oracle import oracle as O
    from oracle import phase2_ref
    from rna_clique_b200.fasta import read_fasta
    from rna_clique_b200.filtered_distance import SampleSimilarity
    from rna_clique_b200.gene_matches_tables import default_extension, get_table_files, read_table
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import simulate, write_dataset
    from rna_clique_b200.transcripts import parse_ids
    _, ts = simulate(3, 150, seed=21, mutation_rate=0.03, indel_rate=0.002)
    dirs = write_dataset(tmp_path / "in", ts)
    od1, od2 = tmp_path / "od1", tmp_path / "od2"
    od1.mkdir(); od2.mkdir()
    sim, pts = rna_clique(dirs, od1, od2, tmp_path / "db", tmp_path / "graph.pkl", tmp_path / "matrix.pkl", 100)
    assert sorted(pts.values()) == ["T0", "T1", "T2"] and all(p.exists() for p in pts)
    assert sim._use_device()                           # reduced on the device-resident tables, od2/ not read back
    ext = default_extension()
    assert sorted(p.name for p in get_table_files(od2)) == [f"T0--T1.{ext}", f"T0--T2.{ext}", f"T1--T2.{ext}"]
    mat = pd.read_pickle(tmp_path / "matrix.pkl")
    labels = sorted(str(p) for p in pts)
    assert list(mat.index) == labels and (np.diag(mat.to_numpy()) == 0).all()
    # oracle: same top-gene files, oracle search, reference-pinned phase 2
    osamples, paths = [], list(pts)
    for p in paths:
        fa = read_fasta(p)
        assert len(fa) == 100
        _, gene, iso = parse_ids(fa.ids)
        osamples.append(O.Sample(fa.ascii, fa.offsets, gene, iso))
    otables = []
    for i in range(3):
        for j in range(i + 1, 3):
            want = O.search_pair(osamples[i], osamples[j]).rows(2)
            got = read_table(od2 / f"T{i}--T{j}.{ext}")
            assert got["ssample"][0] == str(paths[i]) and got["qsample"][0] == str(paths[j])
            for col, k in (("length", 0), ("mismatch", 1), ("gapopen", 2), ("qstart", 3), ("qend", 4), ("sstart", 5),
                           ("send", 6), ("gaps", 8), ("nident", 9), ("qgene", 10), ("qiso", 11), ("sgene", 12), ("siso", 13)):
                assert np.array_equal(got[col].to_numpy().astype(np.int64), want[:, k]), col
            assert np.array_equal(got["reverse"].to_numpy(), want[:, 16].astype(bool))
            assert np.array_equal(got["sstrand"].to_numpy() == "minus", want[:, 15].astype(bool))
            assert (got["evalue"] <= 1e-99).all() and (got["bitscore"] > 300).all()
            otables.append(dict(s=i, q=j, sgene=want[:, 12], qgene=want[:, 10], nident=want[:, 9], length=want[:, 0], gaps=want[:, 8]))
    exp = phase2_ref.phase2(3, otables)
    _, m = phase2_ref.distance_matrix([str(p) for p in paths], otables, exp["sums"])
    assert np.array_equal(mat.to_numpy(), m)
    assert sim.ideal_component_count == exp["n_ideal"] > 50
    with open(tmp_path / "graph.pkl", "rb") as f:
        g = pickle.load(f)
    assert type(g) is nx.Graph and g.number_of_edges() == exp["n_edges"]
    # phase-2-only entry on the files just written
    again = SampleSimilarity.from_filenames(tmp_path / "graph.pkl", sorted(get_table_files(od2)))
    assert np.array_equal(again.get_dissimilarity_df().to_numpy(), m)
    assert sorted(again.valid.itertuples(index=False, name=None)) == sorted(sim.valid.itertuples(index=False, name=None))
    # touching the graph through networkx drops the device fast path; the tables are then read from od2/
    sim2, _ = rna_clique(dirs, od1, od2, tmp_path / "db", tmp_path / "graph.pkl", None, 100, top_matches=2, keep_all=False)
    g2 = sim2.graph
    g2.add_node(("extra", 1)); g2.remove_node(("extra", 1))
    assert not sim2._use_device()
    want2 = [O.search_pair(osamples[i], osamples[j], top_n=2, keep_all=False).rows(2) for i in range(3) for j in range(i + 1, 3)]
    ot2 = [dict(s=i, q=j, sgene=w[:, 12], qgene=w[:, 10], nident=w[:, 9], length=w[:, 0], gaps=w[:, 8])
           for (i, j), w in zip([(0, 1), (0, 2), (1, 2)], want2)]
    e2 = phase2_ref.phase2(3, ot2)
    _, m2 = phase2_ref.distance_matrix([str(p) for p in paths], ot2, e2["sums"])
    assert np.array_equal(sim2.get_dissimilarity_df().to_numpy(), m2)


def test_rna_clique_writes_matrix_h5(tmp_path):
    """The on-disk contract of the metric: rna_clique(..., output_matrix="distance_matrix.h5") (rna_clique.py:175-177)."""
    from rna_clique_b200 import h5fixed
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import simulate, write_dataset
    _, ts = simulate(4, 120, seed=3, mutation_rate=0.02)
    dirs = write_dataset(tmp_path / "in", ts)
    (tmp_path / "od1").mkdir(); (tmp_path / "od2").mkdir()
    sim, pts = rna_clique(dirs, tmp_path / "od1", tmp_path / "od2", tmp_path / "db", tmp_path / "graph.pkl",
                          tmp_path / "distance_matrix.h5", 90, jobs=3)
    m = h5fixed.read_frame(tmp_path / "distance_matrix.h5", "matrix")
    assert m.equals(sim.get_dissimilarity_df()) and m.shape == (4, 4) and list(m.index) == sorted(str(p) for p in pts)
    assert len(list((tmp_path / "od2").iterdir())) == 6


def test_run_analysis_from_yaml(tmp_path):
    """config.yaml -> run_analysis (what the rna-clique command does, rna_clique.py:180-217): output layout,
    matrix file and the finished config written back."""
    from rna_clique_b200.config import RNACliqueConfig, run_analysis
    from rna_clique_b200.simulate import simulate, write_dataset
    _, ts = simulate(3, 120, seed=5, mutation_rate=0.02)
    dirs = write_dataset(tmp_path / "in", ts)
    cfg = RNACliqueConfig(input_dirs=dirs, output_dir=tmp_path / "out", top_genes=80)
    cfg.yaml_save(tmp_path / "config.yaml")
    cfg = RNACliqueConfig.yaml_load(tmp_path / "config.yaml")
    sim, mat = run_analysis(cfg, tmp_path / "out" / "config.yaml")
    assert mat.shape == (3, 3) and (np.diag(mat.to_numpy()) == 0).all() and sim.ideal_component_count > 40
    assert (tmp_path / "out" / "graph.pkl").is_file() and cfg.matrix.is_file() and len(list((tmp_path / "out" / "od2").iterdir())) == 3
    done = RNACliqueConfig.yaml_load(tmp_path / "out" / "config.yaml")
    assert done.finished is not None and sorted(done.path_to_sample.values()) == ["T0", "T1", "T2"]


def test_preloaded_sample_equals_the_one_read_back_from_od1(tmp_path):
    """filtering_step hands the selected records to the device while writing od1/*_top.fasta; the sample
    must be the one that parsing the written file gives (what make_all_dbs did before)."""
    from rna_clique_b200.find_homologs import SampleCache
    from rna_clique_b200.select_top_genes import select_top_and_save
    from rna_clique_b200.simulate import simulate, write_dataset
    from rna_clique_b200.transcripts import default_parser
    _, ts = simulate(2, 300, seed=8, mutation_rate=0.02)
    dirs = write_dataset(tmp_path / "in", ts)
    (tmp_path / "od1").mkdir()
    pre, plain = SampleCache(), SampleCache()
    out, _ = select_top_and_save(tmp_path / "od1", "transcripts.fasta", dirs[0], 120, default_parser, preload=pre)
    assert len(pre._samples) == 1
    a, b = pre.get(out, default_parser), plain.get(out, default_parser)
    assert len(pre._samples) == 1                      # served from the preloaded entry, no second build
    assert (a.n_seq, a.total, a.n_index) == (b.n_seq, b.total, b.n_index) and a.n_seq == 120
    da, db = a.download(), b.download()
    for k in da:
        # index entries are sorted by key only: entries of one key come in the order the appends happened to take
        x, y = (np.sort(da[k]), np.sort(db[k])) if k == "index" else (da[k], db[k])
        assert np.array_equal(x, y), k
    pre.close(); plain.close()


def test_pairwise_strands_of_isoforms(ctx):
    """export_orthologs.blast_pairwise_get_strands on the device search: orientation per isoform pair,
    against the same reduction of the oracle's HSP set."""
    from collections import namedtuple
    from oracle import oracle as O
    from rna_clique_b200.pairwise_strands import pairwise_get_strands
    Rec = namedtuple("Rec", "id seq")
    rng = np.random.default_rng(12)
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    base = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=1500))
    other = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=900))
    recs = [(1, Rec("NODE_1_g5_i1", base.decode())),
            (2, Rec("NODE_2_g5_i2", base[200:1300].translate(comp)[::-1].decode())),     # reverse complement of a part
            (3, Rec("NODE_3_g5_i3", (base[:700] + base[760:]).decode())),                # an exon skipped
            (4, Rec("NODE_4_g5_i4", other.decode()))]                                    # unrelated
    got = list(pairwise_get_strands(recs, ctx=ctx))
    assert got == [((2, "NODE_2_g5_i2"), (1, "NODE_1_g5_i1"), -1),
                   ((3, "NODE_3_g5_i3"), (1, "NODE_1_g5_i1"), 1),
                   ((3, "NODE_3_g5_i3"), (2, "NODE_2_g5_i2"), -1)]
    # the oracle's HSPs through the same reduction
    seqs = [r.seq.encode() for _, r in recs]
    offs = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    s = O.Sample(np.frombuffer(b"".join(seqs), dtype=np.uint8), offs, np.arange(4, dtype=np.int32), np.ones(4, dtype=np.int32))
    hs = O.search_pair(s, s, evalue=1.0).hsps()
    names = [r.id for _, r in recs]
    best = {}
    for row in hs.tolist():
        q, t = names[row[1]], names[row[0]]
        if q > t and (best.get((q, t)) is None or row[7] > best[(q, t)][0]):
            best[(q, t)] = (row[7], row[2])
    assert {k: (-1 if v[1] else 1) for k, v in best.items()} == {(a[1], b[1]): sgn for a, b, sgn in got}
    assert list(pairwise_get_strands(recs[:1], ctx=ctx)) == []


@pytest.mark.parametrize("kind", ["orthologs", "families"])
def test_device_graph_pickle_equals_host_writer_and_networkx(tmp_path, kind):
    """rcq_graph_pickle_tables (node numbering, duplicate rows, adjacency and memo numbers on the GPU) writes the
    bytes of the host writer on the downloaded columns, which are the bytes of pickle.dump(networkx.Graph)."""
    import itertools
    import pickle
    import networkx as nx
    from rna_clique_b200 import _native
    from rna_clique_b200.build_graph import EdgeArrays
    from rna_clique_b200.engine import select_top_genes
    ctx = _native.Context(0)
    try:
        if kind == "orthologs":
            from rna_clique_b200.simulate import simulate
            _, ts = simulate(5, 400, seed=21, mutation_rate=0.03)
            samples = []
            for t in ts:
                offs = t.offsets.numpy()
                keep = select_top_genes(t.coverage, t.gene, 300)
                samples.append(_native.Sample(ctx, t.ascii, offs[:-1][keep], t.gene[keep], t.iso[keep],
                                              lengths=np.diff(offs).astype(np.int32)[keep]))
        else:
            from helpers import make_family_samples
            fam = make_family_samples(seed=4, genes=150, taxa=4)
            samples = [_native.Sample(ctx, a, o, g, i) for a, o, g, i in fam]
        n = len(samples)
        names = [f"/some/dir/T{k}_top.fasta" for k in range(n)]
        pairs = list(itertools.combinations(range(n), 2))
        tables = [_native.search_pair(ctx, samples[i], samples[j], 1e-99, 2 if kind == "families" else 1, True) for i, j in pairs]
        dev = _native.graph_pickle_tables(ctx, tmp_path / "dev.pkl", names, tables, [i for i, _ in pairs], [j for _, j in pairs])
        e = EdgeArrays()
        for (i, j), t in zip(pairs, tables):
            c = t.download()
            if len(c["sgene"]):
                e.add(names[i], names[j], c["sgene"], c["qgene"])
        host = e.write_pickle(tmp_path / "host.pkl")
        assert dev == host and dev["edges"] > 100
        assert (tmp_path / "dev.pkl").read_bytes() == (tmp_path / "host.pkl").read_bytes()
        g = pickle.load(open(tmp_path / "dev.pkl", "rb"))
        ref = e.to_networkx()
        assert type(g) is nx.Graph and list(g.nodes) == list(ref.nodes) and list(g.edges) == list(ref.edges)
        assert all(list(g.adj[x]) == list(ref.adj[x]) for x in ref.nodes)            # neighbours in insertion order
        u, v = next(iter(g.edges))
        assert g._adj[u][v] is g._adj[v][u] and g._adj[u][v] == {}                    # one shared attribute dict per edge
    finally:
        ctx.close()
