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
    """Winnowing guarantee: every 28-base window free of seeding-mask bits (N or DUST) holds at least one
    indexed 16-mer, and no indexed 16-mer touches a masked base."""
    _, raw = make_samples(taxa=1, count=30, seed=3)
    s = _gpu_samples(ctx, raw)[0]
    d = s.download()
    pos = ((d["index"] & np.uint64(0xFFFFFFFF)) >> np.uint64(1)).astype(np.int64)
    assert np.all(np.diff((d["index"] >> np.uint64(32)).astype(np.int64)) >= 0)     # sorted by key
    sel = np.zeros(s.arena, dtype=bool)
    sel[pos] = True
    offsets = raw[0][1]
    words, n_dust = s.seed_mask()
    checked = 0
    for t in range(s.n_seq):
        n = int(offsets[t + 1] - offsets[t])
        o = int(d["off"][t])
        cs = np.concatenate([[0], np.cumsum(sel[o:o + n - 15])])
        win = cs[13:] - cs[:-13]           # windows of 13 consecutive 16-mers
        m = unpack_mask(words, o, n)
        cm = np.concatenate([[0], np.cumsum(m)])
        clean = (cm[28:] - cm[:-28]) == 0  # 28-base windows without a masked base
        assert win[clean].min() >= 1
        checked += int(clean.sum())
        k16 = (cm[16:] - cm[:-16]) > 0     # 16-mers touching a masked base
        assert not (sel[o:o + n - 15] & k16).any()
    assert checked > 0.9 * (offsets[-1] - offsets[0] - 27 * s.n_seq) and n_dust > 0


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
    if tab.stats["band_retries"] == 0:
        # the work unit of bench.py's GCUPS: base comparisons of the gapped stage, counted alike on both sides
        assert tab.stats["dp_cells"] == res.counts["dp_cells"]


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
def test_drifting_band(ctx, length):
    """One base of every 29 deleted in sample 2: exact blocks of 28 (just enough to seed) and a band that
    drifts one diagonal per block, more than a hundred diagonals in all.  The half-warp kernel re-centres its
    48-diagonal register window as the band moves (4000-base transcripts); 6500-base transcripts are longer than
    the staged limit and go to the wide kernel on global rows."""
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
    if length > 4096:
        assert tab.stats["band_retries"] >= n          # every group deferred: transcripts beyond the staged length
    else:
        assert tab.stats["band_retries"] == 0 and tab.stats["dp_cells"] == res.counts["dp_cells"]
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    cols = tab.download()
    got = np.stack([cols[c] for c in _native.TABLE_COLS] + [cols["minus"].astype(np.int32), cols["reverse"].astype(np.int32)], axis=1)
    assert np.array_equal(got, res.rows(2)) and len(tab) == 2 * n
    assert (cols["gaps"] > 100).all()


@pytest.mark.parametrize("unit,reps", [(11, 70), (7, 100), (17, 50)])
def test_tandem_repeats_widen_the_band_beyond_the_register_window(ctx, unit, reps):
    """Unique flanks around a tandem repeat with a few units missing in sample 2: inside the repeat every
    diagonal a multiple of the unit away matches as well, the wave widens past 48 diagonals (the oracle reports
    its widest wave) and the group is finished by the wide kernel; HSPs and table equal the oracle's."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    rng = np.random.default_rng(4)
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    s1, s2 = [], []
    for _ in range(10):
        u = rng.integers(0, 4, size=unit)
        a = np.concatenate([rng.integers(0, 4, size=300), np.tile(u, reps), rng.integers(0, 4, size=300)])
        b = a.copy()
        hit = rng.random(len(b)) < 0.02
        b[hit] = (b[hit] + rng.integers(1, 4, size=int(hit.sum()))) & 3
        b = np.delete(b, np.arange(300 + unit * 10, 300 + unit * 13))
        s1.append(a); s2.append(b)

    def pack(ss):
        offs = np.concatenate([[0], np.cumsum([len(x) for x in ss])]).astype(np.int64)
        return letters[np.concatenate(ss)], offs, np.arange(len(ss), dtype=np.int32), np.ones(len(ss), dtype=np.int32)

    raw = [pack(s1), pack(s2)]
    g = _gpu_samples(ctx, raw)
    o = _oracle_samples(raw)
    O.max_band()
    res = O.search_pair(o[0], o[1])
    assert O.max_band() > 48
    tab = _native.search_pair(ctx, g[0], g[1])
    assert tab.stats["band_retries"] > 0
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    assert np.array_equal(_table_matrix(tab), res.rows(2)) and len(tab) >= 20


def test_dust_mask_equals_the_oracle(ctx):
    """Seeding mask of the sample build (k_dust_*) = not-ACGT | symmetric DUST of the oracle, base for base, on
    transcripts mixing random stretches, homopolymers, di-/tri-nucleotide repeats, AT-rich stretches, N's, very
    short transcripts and one of 15 kb."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    from test_dust_oracle import LET, low_complexity_mix
    rng = np.random.default_rng(11)
    lens = [1, 2, 3, 7, 8, 20, 63, 64, 65, 130, 700, 2050, 15000] + [int(x) for x in rng.integers(30, 1500, size=60)]
    seqs = []
    for k, n in enumerate(lens):
        a = LET[low_complexity_mix(rng, n)] if k % 4 else LET[rng.integers(0, 4, size=n)]
        a = a.copy()
        if k % 5 == 0 and n > 10:
            a[rng.integers(0, n, size=3)] = ord("N")
        seqs.append(a)
    ascii_ = np.concatenate(seqs)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    ids = np.arange(len(lens), dtype=np.int32)
    s = _native.Sample(ctx, ascii_, offsets, ids, np.ones(len(lens), dtype=np.int32))
    words, n_dust = s.seed_mask()
    off = s.download()["off"]
    want = O.Sample(ascii_, offsets, ids, np.ones(len(lens), dtype=np.int32)).dust()
    codes = np_pack(ascii_, offsets)
    total_dust = 0
    for t, n in enumerate(lens):
        got = unpack_mask(words, off[t], n)
        exp = want[offsets[t]:offsets[t + 1]] | (codes[offsets[t]:offsets[t + 1]] > 3)
        assert np.array_equal(got, exp.astype(np.uint8)), f"transcript {t} (length {n})"
        total_dust += int((want[offsets[t]:offsets[t + 1]] & (codes[offsets[t]:offsets[t + 1]] < 4)).sum())
    assert n_dust == total_dust and total_dust > 5000
    # filter off: the seeding mask is the N mask
    was = ctx.set_dust(False)
    try:
        s2 = _native.Sample(ctx, ascii_, offsets, ids, np.ones(len(lens), dtype=np.int32))
        w2, nd2 = s2.seed_mask()
        assert nd2 == 0
        for t, n in enumerate(lens):
            assert np.array_equal(unpack_mask(w2, off[t], n), (codes[offsets[t]:offsets[t + 1]] > 3).astype(np.uint8))
    finally:
        ctx.set_dust(was)


def test_shared_low_complexity_tails_are_masked_not_fatal(ctx):
    """Thousands of transcripts sharing a poly-A tail: without the filter every tail seeds against every other
    one (9 million runs here, quadratic in the sample size: the round-1 library stopped with RCQ_ERR_CAPACITY
    at 50,000 genes).  With DUST the tails take no part in seeding and the table is the oracle's."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    rng = np.random.default_rng(2)
    n, body, tail = 3000, 300, 40
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs = np.concatenate([np.concatenate([letters[rng.integers(0, 4, size=body)], np.full(tail, ord("A"), dtype=np.uint8)])
                           for _ in range(n)])
    offs = np.arange(n + 1, dtype=np.int64) * (body + tail)
    ids = np.arange(n, dtype=np.int32)
    iso = np.ones(n, dtype=np.int32)
    a = _native.Sample(ctx, seqs, offs, ids, iso)
    b = _native.Sample(ctx, seqs.copy(), offs, ids, iso)
    tab = _native.search_pair(ctx, a, b)
    res = O.search_pair(O.Sample(seqs, offs, ids, iso), O.Sample(seqs, offs, ids, iso))
    assert tab.stats["n_seeds"] == res.counts["seeds"] == n
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    assert np.array_equal(_table_matrix(tab), res.rows(2)) and len(tab) == 2 * n
    was = ctx.set_dust(False)
    try:
        m = 300                                       # filter off (blastn -dust no): every tail against every tail
        a2 = _native.Sample(ctx, seqs[:offs[m]], offs[:m + 1], ids[:m], iso[:m])
        t2 = _native.search_pair(ctx, a2, a2)
        assert t2.stats["n_seeds"] >= m * m
        a3 = _native.Sample(ctx, seqs, offs, ids, iso)
        with pytest.raises(RuntimeError, match="low-complexity"):      # and at 3,000 transcripts the lists overflow
            _native.search_pair(ctx, a3, a3)
        a2.close(); a3.close()
    finally:
        ctx.set_dust(was)
    a.close(); b.close()


@pytest.mark.parametrize("seed", [1, 2])
def test_real_assembly_shapes_tails_microsatellites_long_transcripts(ctx, seed):
    """Isoform / paralog families where a third of the transcripts end in a poly-A tail or a (CA)n stretch, plus
    15-kb transcripts: HSPs, seed count and table equal to the oracle's."""
    from helpers import make_family_samples
    from rna_clique_b200 import _native
    from oracle import oracle as O
    raw = make_family_samples(seed=seed, genes=45, tails=30)
    rng = np.random.default_rng(seed)
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)
    long_root = rng.integers(0, 4, size=15000)
    out = []
    for k, (a, offs, gene, iso) in enumerate(raw):
        x = long_root.copy()
        hit = rng.random(len(x)) < 0.02
        x[hit] = (x[hit] + rng.integers(1, 4, size=int(hit.sum()))) & 3
        x = np.delete(x, rng.integers(0, len(x), size=20))
        a = np.concatenate([a, letters[x], np.full(50, ord("A"), dtype=np.uint8)])
        offs = np.concatenate([offs, [offs[-1] + len(x) + 50]]).astype(np.int64)
        out.append((a, offs, np.concatenate([gene, [7 + k]]).astype(np.int32), np.concatenate([iso, [1]]).astype(np.int32)))
    g = _gpu_samples(ctx, out)
    o = _oracle_samples(out)
    assert g[0].seed_mask()[1] > 500
    tab = _native.search_pair(ctx, g[0], g[1])
    res = O.search_pair(o[0], o[1])
    assert tab.stats["n_seeds"] == res.counts["seeds"]
    assert np.array_equal(_native.debug_last_hsps(ctx), res.hsps())
    assert np.array_equal(_table_matrix(tab), res.rows(2)) and len(tab) > 40
    assert (res.rows(2)[:, 0] > 14000).any()                      # the 15-kb pair is in the table


def test_poly_a_on_a_third_of_50000_genes_does_not_overflow(ctx):
    """The failure the round-1 library had on real assemblies: 50,000 genes, 30 % with a 40-base poly-A tail.
    Every gene still finds its ortholog and nothing else."""
    import torch
    from rna_clique_b200 import _native
    n, body, tail = 50_000, 400, 40
    g = torch.Generator().manual_seed(3)
    base = torch.randint(0, 4, (n, body), generator=g, dtype=torch.uint8)
    letters = torch.tensor([65, 67, 71, 84], dtype=torch.uint8)
    with_tail = (torch.arange(n) % 10) < 3

    def sample(mut_seed):
        gg = torch.Generator().manual_seed(mut_seed)
        x = base.clone()
        hit = torch.rand((n, body), generator=gg) < 0.01
        x[hit] = (x[hit] + torch.randint(1, 4, (int(hit.sum()),), generator=gg, dtype=torch.uint8)) & 3
        rows = torch.cat([letters[x.long()], torch.full((n, tail), 65, dtype=torch.uint8)], dim=1)
        lens = torch.where(with_tail, body + tail, body)
        keep = torch.arange(body + tail)[None, :] < lens[:, None]
        return rows[keep].numpy(), np.concatenate([[0], np.cumsum(lens.numpy())]).astype(np.int64)

    ids, iso = np.arange(n, dtype=np.int32), np.ones(n, dtype=np.int32)
    a1, o1 = sample(1)
    a2, o2 = sample(2)
    s1, s2 = _native.Sample(ctx, a1, o1, ids, iso), _native.Sample(ctx, a2, o2, ids, iso)
    tab = _native.search_pair(ctx, s1, s2)
    cols = tab.download()
    assert len(tab) == 2 * n and (cols["qgene"] == cols["sgene"]).all()
    assert tab.stats["n_seeds"] < 20 * n
    s1.close(); s2.close()


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
