"""The reference's own real-BLAST rows (docs/formats.md:260-273, committed as
tests/golden/blast_rows_formats_md.json by tests/golden/make_blast_rows.py) against the scoring and
statistics this repository restates: raw score of +1/-2 with linear gap cost 2.5, lambda/K of the bit
score, BLAST's truncating integer print, pident, the printed e-value.  CPU only; the library is
loaded for its host-side statistics exports (no device call)."""
import json
import math
from pathlib import Path

import numpy as np
import pytest

ROWS = json.loads((Path(__file__).parent / "golden" / "blast_rows_formats_md.json").read_text())["rows"]


def raw_score(r):
    # doubled units: match +2, mismatch -4, gap character -5 (csrc/common.cuh); BLAST reports floor(score2 / 2)
    return (2 * r["nident"] - 4 * r["mismatch"] - 5 * r["gaps"]) >> 1


def test_rows_are_complete_and_consistent():
    assert len(ROWS) == 10
    for r in ROWS:
        assert r["length"] == r["nident"] + r["mismatch"] + r["gaps"]
        q = abs(r["qend"] - r["qstart"]) + 1
        s = abs(r["send"] - r["sstart"]) + 1
        assert q + s == 2 * r["length"] - r["gaps"]
        assert (r["sstrand"] == "minus") == (r["sstart"] > r["send"])
        assert r["gapopen"] <= r["gaps"]


@pytest.mark.parametrize("impl", ["hand", "oracle", "library"])
def test_bitscore_of_every_real_row(impl):
    """bitscore column = trunc((1.28 S - ln 0.46) / ln 2) for S from the integer columns: all ten rows."""
    if impl == "hand":
        bits = lambda s: (1.28 * s - math.log(0.46)) / math.log(2.0)
    elif impl == "oracle":
        from oracle import oracle as O
        bits = O.lib().orc_bitscore
    else:
        from rna_clique_b200 import _native
        bits = _native.lib().rcq_bitscore
    for r in ROWS:
        b = bits(raw_score(r))
        assert b > 99.9 and int(b) == r["bitscore"], (r, b)


def test_printed_columns_match_blast():
    """The frame columns the product derives on the host: bitscore truncated (rounding fails 5 of the 10
    rows), pident to three decimals, e-values below 1e-180 printed as 0.0."""
    from rna_clique_b200.find_homologs import KA_K, KA_LAMBDA, blast_printed_bitscore, blast_printed_evalue
    s = np.array([raw_score(r) for r in ROWS])
    b = (KA_LAMBDA * s - np.log(KA_K)) / np.log(2.0)
    assert blast_printed_bitscore(b).tolist() == [float(r["bitscore"]) for r in ROWS]
    assert int((np.round(b) != blast_printed_bitscore(b)).sum()) == 5
    pid = np.round(100.0 * np.array([r["nident"] for r in ROWS]) / np.array([r["length"] for r in ROWS]), 3)
    assert pid.tolist() == [r["pident"] for r in ROWS]
    # the one non-zero e-value: 8.75e-152 at S = 292 -> effective search space K m' n' = E e^(lambda S)
    r = next(r for r in ROWS if r["evalue"] != 0)
    space = r["evalue"] * math.exp(KA_LAMBDA * raw_score(r)) / KA_K
    assert 1e10 < space < 1e13          # ~2 kb query x ~2e8-base top-gene database
    # every other row is below 1e-180 for any search space up to 1e15 and is printed as 0.0
    for r2 in ROWS:
        if r2 is r:
            continue
        e = KA_K * 1e15 * math.exp(-KA_LAMBDA * raw_score(r2))
        assert e < 1e-180 and r2["evalue"] == 0
    e = np.array([8.7512e-152, 3e-200, 0.0, 1.2345e-100, 0.00012349, 0.012345, 0.56789, 3.14159, 42.4])
    assert blast_printed_evalue(e).tolist() == [8.75e-152, 0.0, 0.0, 1.23e-100, 1.23e-4, 0.012, 0.57, 3.1, 42.0]


def test_length_adjustment_hand_computed():
    """BLAST_ComputeLengthAdjustment for blastn 1/-2 (alpha/lambda = 1.171875, beta = -2, K = 0.46), checked
    against values worked out by hand from the fixed-point definition instead of against the other copy of
    the same code: ell is the largest integer with  alpha/lambda * ln(K (m - ell)(n - N ell)) + beta >= ell."""
    from oracle import oracle as O
    from rna_clique_b200 import _native

    def fixed_point(m, n, N):
        best = 0
        for ell in range(0, int(min(m, n / N))):
            ss = (m - ell) * (n - N * ell)
            if ss <= 0:
                break
            if 1.5 / 1.28 * (math.log(0.46) + math.log(ss)) - 2.0 >= ell:
                best = ell
        return best

    for m, n, N in [(2050, 102_500_000, 50_000), (298, 205_000_000, 90_000), (1578, 4_000_000, 2000), (500, 1_000_000, 400),
                    (15_000, 250_000_000, 120_000), (60, 100_000, 50)]:
        want = fixed_point(m, n, N)
        assert O.lib().orc_length_adjustment(m, n, N) == want, (m, n, N)
        assert _native.lib().rcq_length_adjustment(m, n, N) == want, (m, n, N)
    # docs/formats.md:271: E = 8.75e-152 at S = 292 is reproduced by a 1,578+ base query against a database of
    # the size of a 50,000-gene rnaSPAdes top set (here: m = 1900, n = 2.12e8, N = 105,000 gives 8.7e-152)
    e = _native.lib().rcq_evalue(292, 1900, 212_000_000, 105_000)
    assert 5e-152 < e < 1.5e-151


def test_native_table_statistics_equal_the_numpy_statement():
    """rcq_table_stats (what table_to_frame calls) against find_homologs.table_stats_numpy on rows that reach every
    printing branch: e-values below 1e-180, between 1e-180 and 0.0009, and up to tens; bit scores on both sides of
    99.9; both directions with different database sizes."""
    import numpy as np
    from rna_clique_b200 import find_homologs as fh
    rng = np.random.default_rng(1)
    n = 60000
    score = np.concatenate([rng.integers(8, 60, size=n // 3), rng.integers(60, 400, size=n // 3), rng.integers(400, 4200, size=n - 2 * (n // 3))]).astype(np.int32)
    rng.shuffle(score)
    length = rng.integers(28, 4200, size=n).astype(np.int32)
    cols = dict(score=score, qlen=np.repeat(rng.integers(30, 9000, size=n // 4), 4).astype(np.int32), length=length,
                nident=(length - rng.integers(0, 28, size=n)).astype(np.int32), reverse=(rng.integers(0, 2, size=n)).astype(np.uint8))
    for k in ("mismatch", "gapopen", "qstart", "qend", "sstart", "send", "gaps", "qgene", "qiso", "sgene", "siso"):
        cols[k] = rng.integers(0, 300, size=n).astype(np.int32)
    cols["minus"] = rng.integers(0, 2, size=n).astype(np.uint8)
    for db in ((102500000, 50000, 99000000, 48000), (12000, 7, 3000000, 900)):
        want = fh.table_stats_numpy(cols, *db)
        df = fh.table_to_frame(cols, *db)
        assert np.array_equal(df["pident"].to_numpy(), want[0])
        assert np.array_equal(df["evalue"].to_numpy(), want[1])
        assert np.array_equal(df["bitscore"].to_numpy(), want[2])
        e = want[1]
        assert (e == 0).any() and ((e > 0) & (e < 0.0009)).any() and (e > 0.0009).any()
        assert (want[2] < 99.9).any() and (want[2] > 99.9).any()
