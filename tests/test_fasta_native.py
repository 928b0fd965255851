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
