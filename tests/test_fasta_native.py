"""Native FASTA ingest and rnaSPAdes id decoding (host code of the library) against the numpy / re
statements of the same contracts (select_top_genes.py:110-127, transcripts.py:8,69-126)."""
import numpy as np
import pytest

from rna_clique_b200.fasta import read_fasta, read_fasta_numpy
from rna_clique_b200.transcripts import TranscriptIDParseError, default_parser, parse_fasta_ids, parse_ids


def _same(a, b):
    assert a.headers == b.headers
    assert np.array_equal(a.ascii, b.ascii) and np.array_equal(a.offsets, b.offsets)


@pytest.mark.parametrize("text", [
    b">NODE_1_length_20_cov_3.5_g0_i0 some description\nACGTACGT\nACG\n>NODE_2_length_5_cov_10.25_g1_i2\nTTTTT\n",
    b">a_cov_1.0_g1_i1\r\nAC GT\r\nNN\r\n>b_cov_2.0_g2_i1\r\n\r\n>c_cov_3.0_g3_i1\r\nA",          # CRLF, spaces, empty record, no final newline
    b"ACGT\n>late_cov_1.5_g4_i1\nGG\n",                                                          # letters before the first header
    b">only_cov_0.5_g9_i9\n",
    b"\n\n>x_cov_1.5_g1_i1\n\nAC\n\nGT\n",
])
def test_read_fasta_native_equals_numpy(tmp_path, text):
    p = tmp_path / "t.fasta"
    p.write_bytes(text)
    _same(read_fasta(p), read_fasta_numpy(p))


def test_read_fasta_random_files(tmp_path):
    rng = np.random.default_rng(3)
    for trial in range(20):
        parts = []
        for r in range(int(rng.integers(0, 30))):
            parts.append(b">NODE_%d_length_9_cov_%d.%d_g%d_i%d extra\n" % (r, rng.integers(0, 999), rng.integers(0, 99), rng.integers(0, 50), rng.integers(1, 4)))
            seq = bytes(rng.choice(np.frombuffer(b"ACGTNacgt", dtype=np.uint8), size=int(rng.integers(0, 200))))
            w = int(rng.integers(1, 80))
            parts.append(b"\n".join(seq[i:i + w] for i in range(0, len(seq), w)) + (b"\n" if rng.integers(0, 2) else b"\r\n"))
        p = tmp_path / f"r{trial}.fasta"
        p.write_bytes(b"".join(parts))
        a, b = read_fasta(p), read_fasta_numpy(p)
        _same(a, b)
        if len(a):
            for x, y in zip(parse_fasta_ids(a), parse_ids(a.ids)):
                assert np.array_equal(x, y)


def test_default_pattern_semantics(tmp_path):
    ids = ["NODE_1_length_15273_cov_210.5_g0_i0", "x_cov_1.5_g1_i1_cov_2.25_g7_i3", "pre_cov_3.0_g12_i4trailing",
           "cov_0.125_g5_i6", "a_cov_7.5_g1_i1_cov_9_g2_i2"]     # the last one: second cov_ has no decimal point
    p = tmp_path / "ids.fasta"
    p.write_text("".join(f">{i} desc\nACGT\n" for i in ids))
    fa = read_fasta(p)
    cov, gene, iso = parse_fasta_ids(fa)
    want = [default_parser(i) for i in ids]
    assert cov.tolist() == [w.coverage for w in want] and gene.tolist() == [w.gene for w in want] and iso.tolist() == [w.isoform for w in want]
    assert (gene.tolist(), iso.tolist()) == ([0, 7, 12, 5, 1], [0, 3, 4, 6, 1])
    bad = tmp_path / "bad.fasta"
    bad.write_text(">ok_cov_1.5_g1_i1\nAC\n>NODE_7_cov_12_g1_i1\nAC\n")          # integer coverage: no match upstream either
    with pytest.raises(TranscriptIDParseError, match="NODE_7_cov_12_g1_i1"):
        parse_fasta_ids(read_fasta(bad))


def test_write_fasta_native_equals_python(tmp_path):
    from rna_clique_b200.fasta import write_fasta, write_fasta_python
    rng = np.random.default_rng(1)
    lens = [0, 1, 59, 60, 61, 120, 121, 500]
    ascii_ = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=sum(lens) + 7)
    starts = np.concatenate([[3], 3 + np.cumsum(lens)[:-1]])
    headers = [f"NODE_{k}_cov_1.5_g{k}_i1 d" for k in range(len(lens))]
    for w in (60, 7):
        write_fasta(tmp_path / "a.fa", headers, ascii_, starts, lens, w)
        write_fasta_python(tmp_path / "b.fa", headers, ascii_, starts, lens, w)
        assert (tmp_path / "a.fa").read_bytes() == (tmp_path / "b.fa").read_bytes()
    back = read_fasta(tmp_path / "a.fa")
    assert back.headers == headers and np.array_equal(np.diff(back.offsets), lens)


def test_lean_parse_equals_the_full_parse(tmp_path):
    """_parse_and_select_lean (mapped file, header spans, letters of the selected records only) hands on the same
    records and writes the same od1 file as the full parse - CRLF line ends, spaces inside sequence lines, ragged
    line lengths, a last record without a line end, isoforms of one gene, a custom buffer allocator."""
    from rna_clique_b200 import select_top_genes as stg
    rng = np.random.default_rng(3)
    lines = []
    for g in range(40):
        for i in range(1 + g % 3):
            n = int(rng.integers(1, 400))
            seq = "".join(rng.choice(list("ACGTN"), size=n))
            w = int(rng.integers(7, 90))
            body = [seq[k:k + w] for k in range(0, n, w)]
            if g % 5 == 0:
                body = [b[:3] + " " + b[3:] for b in body]
            eol = "\r\n" if g % 4 == 1 else "\n"
            lines.append(f">NODE_{g}_length_{n}_cov_{rng.uniform(1, 900):.4f}_g{g}_i{i} extra words" + eol + eol.join(body) + eol)
    text = "".join(lines)
    text = text[:-1]                                             # no line end after the last record
    (tmp_path / "a").mkdir()
    (tmp_path / "a" / "transcripts.fasta").write_text(text, newline="")
    for sub in ("full", "lean"):
        (tmp_path / sub).mkdir()
    asked = []

    def alloc(nbytes):
        asked.append(nbytes)
        return np.empty(nbytes + 100, dtype=np.uint8)

    for top in (0, 1, 7, 40, 100):
        o1, s1, r1, w1 = stg._parse_and_select(tmp_path / "full", "transcripts.fasta", tmp_path / "a", top, default_parser)
        o2, s2, r2, w2 = stg._parse_and_select_lean(tmp_path / "lean", "transcripts.fasta", tmp_path / "a", top, default_parser, alloc)
        w1(); w2()
        assert s1 == s2 and o1.name == o2.name
        assert o1.read_bytes() == o2.read_bytes()
        for k in (3, 4, 5):
            assert np.array_equal(r1[k], r2[k])
        for a0, n, b0 in zip(r1[2], r1[5], r2[2]):
            assert np.array_equal(r1[1][a0:a0 + n], r2[1][b0:b0 + n])
    assert asked                                                 # the allocator was used
    # any other parser takes the full path
    import re
    custom = lambda s: default_parser(s)
    o3, _, r3, _ = stg._parse_and_select_lean(tmp_path / "lean", "transcripts.fasta", tmp_path / "a", 7, custom)
    assert np.array_equal(r3[3], stg._parse_and_select(tmp_path / "full", "transcripts.fasta", tmp_path / "a", 7, custom)[2][3])
    # an id the pattern does not match raises like the reference
    (tmp_path / "b").mkdir()
    (tmp_path / "b" / "transcripts.fasta").write_text(">NODE_1_cov_3.5_g1_i1\nACGT\n>nonsense\nACGT\n")
    with pytest.raises(TranscriptIDParseError):
        stg._parse_and_select_lean(tmp_path / "lean", "transcripts.fasta", tmp_path / "b", 1, default_parser)


def test_native_coverage_values_equal_python_float():
    """rcq_parse_spades_ids reads short decimals (at most 15 digits) as integer / power of ten - one correctly rounded
    division - and longer ones with strtod: every value must be the double float() gives."""
    from rna_clique_b200 import _native
    rng = np.random.default_rng(5)
    strs = []
    for _ in range(20000):
        d1, d2 = int(rng.integers(1, 10)), int(rng.integers(1, 12))
        strs.append("".join(rng.choice(list("0123456789"), size=d1)) + "." + "".join(rng.choice(list("0123456789"), size=d2)))
    strs += ["0.1", "0.3", "123456789.123456", "99999999.9999999", "1.000000000000001", "12345678901234567.123",
             "0.000000000000001", "2.675", "1.005", "0.29999999999999999", "0.0"]
    text = "".join(f">NODE_{i}_length_10_cov_{s}_g{i % 1000}_i{1 + i % 3}\nACGT\n" for i, s in enumerate(strs)).encode()
    raw = np.frombuffer(text, dtype=np.uint8)
    hs, he = _native.fasta_headers(raw)
    cov, gene, iso, bad = _native.parse_spades_ids(raw, hs, he)
    assert bad == -1
    assert np.array_equal(cov, np.array([float(s) for s in strs]))
    assert np.array_equal(gene, np.arange(len(strs)) % 1000) and np.array_equal(iso, 1 + np.arange(len(strs)) % 3)
