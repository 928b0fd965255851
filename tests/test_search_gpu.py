"""GPU parity: pack / index / per-pair search through the C ABI against the CPU oracle."""
import numpy as np
import pytest

from helpers import make_samples, np_pack, unpack_arena, unpack_mask

pytestmark = pytest.mark.gpu


def _gpu_samples(ctx, raw):
    from rna_clique_b200 import _native
    return [_native.Sample(ctx, *r) for r in raw]


def _oracle_samples(raw):
    from oracle import oracle as O
    return [O.Sample(*r) for r in raw]


def test_pack_revcomp_mask(ctx):
    rng = np.random.default_rng(5)
    lens = [1, 15, 16, 17, 31, 32, 33, 100, 511, 512, 513, 2050, 0, 64]
    # includes every letter that differs from a valid one in a single bit group (E P Q R S V W, @ ` [ {)
    letters = np.frombuffer(b"ACGTacgtNnRYUu-EePpQqSsVvWw@`[{\x00\xff\x14\x95", dtype=np.uint8)
    ascii_ = letters[rng.integers(0, len(letters), size=sum(lens))]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    from rna_clique_b200 import _native
    s = _native.Sample(ctx, ascii_, offsets, np.arange(len(lens), dtype=np.int32), np.ones(len(lens), dtype=np.int32))
    d = s.download()
    codes = np_pack(ascii_, offsets)
    assert s.n_masked == int((codes > 3).sum())
    for t, n in enumerate(lens):
        c = codes[offsets[t]:offsets[t + 1]]
        got = unpack_arena(d["fwd"], d["off"][t], n)
        gm = unpack_mask(d["mask"], d["off"][t], n)
        assert np.array_equal(gm, (c > 3).astype(np.uint8))
        assert np.array_equal(got[c < 4], c[c < 4])
        rc = unpack_arena(d["rc"], d["off"][t], n)
        exp = (3 - c[::-1]).astype(np.uint8)
        ok = c[::-1] < 4
        assert np.array_equal(rc[ok], exp[ok])
        assert d["off"][t] % 32 == 0


def test_index_covers_every_28mer_window(ctx):
    """Winnowing guarantee: every N-free 28-base window holds at least one indexed 16-mer."""
    _, raw = make_samples(taxa=1, count=30, seed=3)
    s = _gpu_samples(ctx, raw)[0]
    d = s.download()
    pos = ((d["index"] & np.uint64(0xFFFFFFFF)) >> np.uint64(1)).astype(np.int64)
    assert np.all(np.diff((d["index"] >> np.uint64(32)).astype(np.int64)) >= 0)     # sorted by key
    sel = np.zeros(s.arena, dtype=bool)
    sel[pos] = True
    offsets = raw[0][1]
    for t in range(s.n_seq):
        n = int(offsets[t + 1] - offsets[t])
        o = int(d["off"][t])
        cs = np.concatenate([[0], np.cumsum(sel[o:o + n - 15])])
        win = cs[13:] - cs[:-13]           # windows of 13 consecutive 16-mers
        assert win.min() >= 1


@pytest.mark.parametrize("kw", [
    dict(taxa=2, count=150, seed=1),
    dict(taxa=2, count=60, seed=2, mutation_rate=0.08),
    dict(taxa=2, count=60, seed=3, mutation_rate=0.03, indel_rate=0.004),
    dict(taxa=2, count=120, seed=4, mutation_rate=0.15, indel_rate=0.01),
    dict(taxa=2, count=80, seed=5, mutation_rate=0.05, indel_rate=0.03),
    dict(taxa=2, count=40, seed=6, mutation_rate=0.02, indel_rate=0.002, len_loc=9000, len_n=4000),
])
def test_search_pair_matches_oracle(ctx, kw):
    from rna_clique_b200 import _native
    from oracle import oracle as O
    _, raw = make_samples(**kw)
    g = _gpu_samples(ctx, raw)
    o = _oracle_samples(raw)
    tab = _native.search_pair(ctx, g[0], g[1])
    res = O.search_pair(o[0], o[1])
    hs = _native.debug_last_hsps(ctx)
    assert tab.stats["n_seeds"] == res.counts["seeds"]
    assert np.array_equal(hs, res.hsps())
    cols = tab.download()
    got = np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)
    assert np.array_equal(got, res.rows(2))
    assert tab.stats["n_fwd"] == res.counts["fwd"] and tab.stats["n_bwd"] == res.counts["bwd"]
    assert len(tab) > 0


def test_search_minus_strand_and_masked_bases(ctx):
    """Half of sample 2's transcripts are reverse-complemented; a few bases become N."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    _, raw = make_samples(taxa=2, count=80, seed=7, mutation_rate=0.02)
    a2, off2, gene2, iso2 = raw[1]
    a2 = a2.copy()
    comp = np.zeros(256, dtype=np.uint8)
    for x, y in zip(b"ACGT", b"TGCA"):
        comp[x] = y
    for t in range(0, len(gene2), 2):
        seg = a2[off2[t]:off2[t + 1]]
        a2[off2[t]:off2[t + 1]] = comp[seg[::-1]]
    rng = np.random.default_rng(0)
    a2[rng.integers(0, len(a2), size=40)] = ord("N")
    raw = [raw[0], (a2, off2, gene2, iso2)]
    g = _gpu_samples(ctx, raw)
    o = _oracle_samples(raw)
    tab = _native.search_pair(ctx, g[0], g[1])
    res = O.search_pair(o[0], o[1])
    hs = _native.debug_last_hsps(ctx)
    assert np.array_equal(hs, res.hsps())
    assert hs[:, 2].sum() > 0
    cols = tab.download()
    got = np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)
    assert np.array_equal(got, res.rows(2))


def test_sample_export_import_round_trip(ctx):
    """The flat device buffer a rank broadcasts (rcq_sample_export / rcq_sample_import) rebuilds a sample
    that searches exactly like the original."""
    import torch
    from rna_clique_b200 import _native
    _, raw = make_samples(taxa=2, count=120, seed=9, mutation_rate=0.03, indel_rate=0.002)
    g = _gpu_samples(ctx, raw)
    meta = g[1].meta()
    buf = torch.empty(int(meta[7]), dtype=torch.uint8, device="cuda:0")
    g[1].export_flat(buf.data_ptr())
    torch.cuda.synchronize()
    clone = _native.Sample.import_flat(ctx, meta, buf.data_ptr(), "clone")
    assert (clone.n_seq, clone.total, clone.n_index) == (g[1].n_seq, g[1].total, g[1].n_index)
    a = _native.search_pair(ctx, g[0], g[1])
    b = _native.search_pair(ctx, g[0], clone)
    ca, cb = a.download(), b.download()
    assert len(a) > 0 and all(np.array_equal(ca[k], cb[k]) for k in ca)
    d0, d1 = g[1].download(), clone.download()
    assert all(np.array_equal(d0[k], d1[k]) for k in d0)


@pytest.mark.parametrize("length", [4000, 6500])
def test_drifting_band_uses_the_wide_fallback(ctx, length):
    """One base of every 29 deleted in sample 2: exact blocks of 28 (just enough to seed) and a band that
    drifts one diagonal per block, far beyond the 124-diagonal shared-memory rows — the extension
    re-runs on the pooled global rows (staged strands for 4000-base transcripts, arena reads for 6500)."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    rng = np.random.default_rng(length)
    n = 12
    seqs = [rng.integers(0, 4, size=length + int(rng.integers(0, 50))) for _ in range(n)]
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    a1 = np.concatenate([letters[s] for s in seqs])
    off1 = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    cut = [np.delete(s, np.arange(28, len(s), 29)) for s in seqs]
    a2 = np.concatenate([letters[s] for s in cut])
    off2 = np.concatenate([[0], np.cumsum([len(s) for s in cut])]).astype(np.int64)
    gene, iso = np.arange(n, dtype=np.int32), np.ones(n, dtype=np.int32)
    raw = [(a1, off1, gene, iso), (a2, off2, gene, iso)]
    g = _gpu_samples(ctx, raw)
    o = _oracle_samples(raw)
    tab = _native.search_pair(ctx, g[0], g[1])
    res = O.search_pair(o[0], o[1])
    assert tab.stats["band_retries"] > 0
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    cols = tab.download()
    got = np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)
    assert np.array_equal(got, res.rows(2)) and len(tab) == 2 * n
    assert (cols["gaps"] > 100).all()


def test_shared_low_complexity_tails_fail_with_a_clear_message(ctx):
    """No DUST filter yet: thousands of transcripts sharing a poly-A tail seed against each other
    quadratically.  The search must stop with RCQ_ERR_CAPACITY and say why, not exhaust the device."""
    from rna_clique_b200 import _native
    rng = np.random.default_rng(2)
    n, body, tail = 5000, 60, 40
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs = np.concatenate([np.concatenate([letters[rng.integers(0, 4, size=body)], np.full(tail, ord("A"), dtype=np.uint8)])
                           for _ in range(n)])
    offs = np.arange(n + 1, dtype=np.int64) * (body + tail)
    ids = np.arange(n, dtype=np.int32)
    a = _native.Sample(ctx, seqs, offs, ids, np.ones(n, dtype=np.int32))
    b = _native.Sample(ctx, seqs[::1].copy(), offs, ids, np.ones(n, dtype=np.int32))
    with pytest.raises(RuntimeError, match="low-complexity"):
        _native.search_pair(ctx, a, b)
    a.close(); b.close()


def _table_matrix(tab):
    from rna_clique_b200 import _native
    cols = tab.download()
    return np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)


@pytest.mark.parametrize("top_n", [1, 2, 3])
@pytest.mark.parametrize("keep_all", [True, False])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_selection_on_isoforms_paralogs_and_sparse_ids(ctx, seed, keep_all, top_n):
    """The table assembly where it has decisions to take (helpers.make_family_samples): several isoforms per
    gene, paralog families with exact copies (equal scores across subject genes), copies present in one
    sample only (non-reciprocal best hits), chimeras, sparse sample-specific gene ids — for keep="all" and
    keep="first" (k_sel_first) and for top_n 1, 2, 3 (k_sel_topn_*), every table equal to the oracle's, whose
    assembly is pinned to the reference's HomologFinder.get_match_table (tests/golden/match_table_*.json)."""
    from helpers import make_family_samples
    from rna_clique_b200 import _native
    from oracle import oracle as O
    raw = make_family_samples(seed=seed)
    g = _gpu_samples(ctx, raw)
    o = _oracle_samples(raw)
    tab = _native.search_pair(ctx, g[0], g[1], top_n=top_n, keep_all=keep_all)
    res = O.search_pair(o[0], o[1], top_n=top_n, keep_all=keep_all)
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    got, want = _table_matrix(tab), res.rows(2)
    assert np.array_equal(got, want)
    assert tab.stats["n_fwd"] == res.counts["fwd"] and tab.stats["n_bwd"] == res.counts["bwd"]
    # the inputs do what they were built for
    q, s = want[:, 10], want[:, 12]
    assert len(want) > 40 and (q != s).all()                         # ids differ between the samples
    pairs = np.unique(np.stack([q, s], axis=1), axis=0)
    if keep_all:
        assert len(np.unique(pairs[:, 0])) < len(pairs)              # a query gene with several subject genes
    else:
        assert len(np.unique(q)) == len(want)                        # keep="first": one row per query gene
    assert len(np.unique(want[:, 11])) > 1                           # several isoform numbers in the rows
    if top_n > 1:
        assert res.counts["fwd"] > O.search_pair(o[0], o[1]).counts["fwd"]


def test_selection_role_swap_on_family_data(ctx):
    """Swapping the two samples swaps q and s of every reciprocal pair: the (gene, gene) edge set is the same."""
    from helpers import make_family_samples
    from rna_clique_b200 import _native
    raw = make_family_samples(seed=5, genes=40)
    g = _gpu_samples(ctx, raw)
    a = _table_matrix(_native.search_pair(ctx, g[0], g[1]))
    b = _table_matrix(_native.search_pair(ctx, g[1], g[0]))
    ea = {(int(r[12]), int(r[10])) for r in a}      # (gene of sample 0, gene of sample 1)
    eb = {(int(r[10]), int(r[12])) for r in b}
    assert len(ea) > 20 and len(ea ^ eb) <= max(1, len(ea) // 50)
