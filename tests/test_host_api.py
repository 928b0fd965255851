"""CPU: host-side mirror of the reference interface (id parsing, FASTA, top genes, tables, graph)."""
import json
import pickle
from fractions import Fraction
from pathlib import Path

import networkx as nx
import numpy as np
import pandas as pd
import pytest

GOLD = Path(__file__).resolve().parent / "golden"


def test_top_genes_against_reference_fixture():
    """tests/golden/top_genes.json was produced by the reference's TopGeneSelector."""
    from rna_clique_b200.select_top_genes import top_gene_mask
    from rna_clique_b200.transcripts import parse_ids
    d = json.loads((GOLD / "top_genes.json").read_text())
    cov, gene, _ = parse_ids(d["ids"])
    for c in d["cases"]:
        k = top_gene_mask(cov, gene, c["top"])
        assert [d["ids"][i] for i in k] == c["kept_ids"]
        assert sorted(set(gene[k].tolist())) == c["genes"]


def test_transcript_id_parser():
    import re
    from rna_clique_b200.transcripts import TranscriptID, TranscriptIDParseError, default_parser
    t = default_parser("NODE_1_length_15273_cov_210.5_g7_i2")
    assert t == (210.5, 7, 2) and isinstance(t.coverage, float) and isinstance(t.gene, int)
    with pytest.raises(TranscriptIDParseError):
        default_parser("NODE_1_length_15273_cov_210_g7_i2")       # coverage needs a decimal point
    trinity = TranscriptID.parser_from_re(re.compile(r"DN(?P<gene>\d+)_c\d+_g\d+_i(?P<isoform>\d+) cov=(?P<coverage>[\d.]+)"))
    assert trinity("TRINITY_DN12_c0_g1_i3 cov=4.25") == (4.25, 12, 3)
    mixed = TranscriptID.parser_from_re(re.compile(r"c([\d.]+)_(?P<isoform>\d+)_(\d+)"))
    assert mixed("c1.5_9_4") == (1.5, 4, 9)


def test_fasta_roundtrip_and_select_top_and_save(tmp_path):
    from rna_clique_b200.fasta import read_fasta
    from rna_clique_b200.select_top_genes import select_top_and_save
    d = tmp_path / "S1"
    d.mkdir()
    recs = [("NODE_1_cov_5.0_g1_i1", "ACGT" * 40), ("NODE_2_cov_9.0_g2_i1 extra words", "GG" * 31),
            ("NODE_3_cov_1.0_g3_i1", "T"), ("NODE_4_cov_7.5_g1_i2", "CA" * 100)]
    (d / "transcripts.fasta").write_text("".join(f">{h}\n" + "\n".join(s[i:i + 70] for i in range(0, len(s), 70)) + "\n" for h, s in recs))
    fa = read_fasta(d / "transcripts.fasta")
    assert fa.headers == [h for h, _ in recs] and fa.ids[1] == "NODE_2_cov_9.0_g2_i1"
    assert [fa.ascii[a:b].tobytes().decode() for a, b in zip(fa.offsets[:-1], fa.offsets[1:])] == [s for _, s in recs]
    out, name = select_top_and_save(tmp_path, "transcripts.fasta", d, 2)
    assert name == "S1" and out == tmp_path / "S1_top.fasta"
    top = read_fasta(out)
    assert top.headers == [recs[0][0], recs[1][0], recs[3][0]]      # genes 2 and 1 (both isoforms), file order
    assert max(len(l) for l in out.read_text().splitlines()) == 60


def test_tables_io_and_paths(tmp_path):
    from rna_clique_b200.find_all_pairs import make_output_path
    from rna_clique_b200.gene_matches_tables import get_table_files, read_table, write_table
    from rna_clique_b200.path_to_sample import PathToSampleError, path_to_sample
    df = pd.DataFrame(dict(qgene=[1, 2], sgene=[1, 2], nident=[10, 20]))
    write_table(df, tmp_path / "a--b.pkl")
    assert read_table(tmp_path / "a--b.pkl").equals(df) and len(read_table(tmp_path / "a--b.pkl", head=1)) == 1
    with pytest.raises(ValueError):
        write_table(df, tmp_path / "a--b.txt")
    assert [p.name for p in get_table_files(tmp_path)] == ["a--b.pkl"]
    assert path_to_sample("od1/SRR1_top.fasta") == "SRR1"
    with pytest.raises(PathToSampleError):
        path_to_sample("od1/SRR1.fasta")
    assert make_output_path(Path("od2"), "x/A_top.fasta", "x/B_top.fasta", path_to_sample, "h5") == Path("od2/A--B.h5")


def test_build_graph_is_lazy_and_pickles_as_networkx():
    from rna_clique_b200.build_graph import build_graph
    t1 = pd.DataFrame(dict(ssample="a", qsample="b", sgene=[1, 2, 2], qgene=[1, 2, 2]))
    t2 = pd.DataFrame(dict(ssample="a", qsample="c", sgene=[1], qgene=[5]))
    g = build_graph(iter([t1, t2]))
    h = pickle.loads(pickle.dumps(g, pickle.HIGHEST_PROTOCOL))
    assert type(h) is nx.Graph
    assert sorted(h.edges()) == [(("a", 1), ("b", 1)), (("a", 1), ("c", 5)), (("a", 2), ("b", 2))]
    assert g.number_of_nodes() == 5 and ("c", 5) in g


def test_pair_dict_and_schedule():
    from rna_clique_b200.engine import schedule_pairs
    from rna_clique_b200.filtered_distance import PairDict
    d = PairDict()
    d[("b", "a")] = Fraction(1, 3)
    d[("a", "a")] = 1
    assert d[["a", "b"]] == d[frozenset(("a", "b"))] == Fraction(1, 3) and d[["a", "a"]] == 1
    assert d.key_elements() == {"a", "b"}
    pairs, owner = schedule_pairs(6, [100, 90, 80, 70, 60, 50], 4)
    assert len(pairs) == 15 and set(owner) == {0, 1, 2, 3}
    load = [sum(100 - 10 * i + 100 - 10 * j for (i, j), o in zip(pairs, owner) if o == r) for r in range(4)]
    assert max(load) - min(load) <= 200


def test_priority_pool_serves_urgent_tasks_first_and_hands_on_errors():
    """filtering_step._PriorityPool: lower priority numbers first, FIFO inside a priority, exceptions through the future."""
    import threading
    from rna_clique_b200.filtering_step import _PriorityPool
    pool = _PriorityPool(1)
    gate, order = threading.Event(), []
    first = pool.submit(0, gate.wait)                     # occupies the only worker while the others queue up
    futs = [pool.submit(p, order.append, (p, k)) for k, p in enumerate([2, 2, 1, 0, 1, 2, 0])]
    boom = pool.submit(1, lambda: 1 / 0)
    gate.set()
    first.result(timeout=10)
    for f in futs:
        f.result(timeout=10)
    assert order == [(0, 3), (0, 6), (1, 2), (1, 4), (2, 0), (2, 1), (2, 5)]
    with pytest.raises(ZeroDivisionError):
        boom.result(timeout=10)
    pool.shutdown()


def test_pinned_pool_reuses_buffers_and_falls_back_to_numpy():
    """filtering_step._PinnedPool: a returned buffer is handed out again; beyond the limit (or without a CUDA device,
    where page-locking fails) the caller gets ordinary memory."""
    from rna_clique_b200.filtering_step import _PinnedPool
    pool = _PinnedPool(limit=64 << 20)
    a = pool.get(1000)
    assert a.dtype == np.uint8 and len(a) >= 1000
    addr = a.ctypes.data
    pool.put(a[:10])                                      # a view that starts at the first byte gives it back
    b = pool.get(500)
    if pool.held:                                         # page-locked memory was available: same buffer again
        assert b.ctypes.data == addr
    big = pool.get(128 << 20)                             # over the limit: plain numpy memory
    assert len(big) >= 128 << 20 and big.ctypes.data not in pool.out
    pool.put(big)
    pool.clear()


def _fake_table(seed, n):
    from rna_clique_b200 import _native
    rng = np.random.default_rng(seed)
    block = rng.integers(1, 2000, size=(len(_native.TABLE_COLS), max(n, 0))).astype(np.int32)
    cols = {c: block[k] for k, c in enumerate(_native.TABLE_COLS)}
    cols["length"][:] = rng.integers(200, 2100, size=n)
    cols["nident"][:] = cols["length"] - rng.integers(0, 30, size=n)
    cols["score"][:] = rng.integers(40, 4000, size=n)
    cols["minus"] = rng.integers(0, 2, size=n).astype(np.uint8)
    cols["reverse"] = (np.arange(n) % 2).astype(np.uint8)
    return cols


def test_table_writer_processes_write_the_files_the_threads_write(tmp_path):
    """filtering_step._TableWriters: od2 files written by the helper processes load as the frames _frame_and_save
    writes in-process; a helper that dies leaves its task to the feeding thread."""
    import pandas as pd
    from rna_clique_b200.filtering_step import _TableWriters, _frame_and_save
    w = _TableWriters(2)
    if not w:
        pytest.skip("helper processes cannot be started here")
    try:
        db = (1_000_000, 500, 900_000, 480)
        tabs = [_fake_table(k, n) for k, n in enumerate([5000, 1, 0, 12000, 300])]
        futs = [w.submit(t, db, "/x/A_top.fasta", "/x/B_top.fasta", tmp_path / f"p{k}.pkl") for k, t in enumerate(tabs)]
        for f in futs:
            f.result(timeout=120)
        for k, t in enumerate(tabs):
            _frame_and_save(t, db, "/x/A_top.fasta", "/x/B_top.fasta", tmp_path / f"t{k}.pkl")
            a, b = pd.read_pickle(tmp_path / f"p{k}.pkl"), pd.read_pickle(tmp_path / f"t{k}.pkl")
            assert a.equals(b) and list(a.dtypes) == list(b.dtypes) and list(a.columns) == list(b.columns)
        assert not list(Path(w.dir).iterdir())                   # the column files are gone
        for p in w.procs:                                        # helpers gone: the feeding threads take over
            p.kill()
            p.wait()
        futs = [w.submit(tabs[0], db, "/x/A_top.fasta", "/x/B_top.fasta", tmp_path / f"q{k}.pkl") for k in range(4)]
        for f in futs:
            f.result(timeout=120)
        assert pd.read_pickle(tmp_path / "q3.pkl").equals(pd.read_pickle(tmp_path / "t0.pkl"))
    finally:
        w.close()
