"""CPU: the oracle's two gapped-extension restatements against each other and against hand cases.
The greedy wave form (canonical; megablast's own algorithm) and the anti-diagonal X-drop DP must
agree on score and end point; only the split of a tie between two mismatches and an
insertion+deletion pair may differ."""
import numpy as np
import pytest

from helpers import make_samples


def test_hand_cases():
    from oracle import oracle as O
    x, y = "ACGTTGCAAGGCTTAGCATCGATCGGATCTAGCTAGGATC", "TTATTCAGCATTCGAGGCTAGCTTTAGGCAGCATCGACTA"
    assert O.extend((x + y).encode(), (x + y).encode()) == dict(score2=160, qi=80, sj=80, mismatch=0, gaps=0, gapopen=0)
    r = O.extend((x + "A" + y).encode(), (x + "G" + y).encode())
    assert (r["score2"], r["mismatch"], r["gaps"]) == (2 * 80 - 4, 1, 0)
    r = O.extend((x + "ACG" + y).encode(), (x + y).encode())          # 3-base insertion in q: one opening
    assert (r["score2"], r["qi"], r["sj"], r["mismatch"], r["gaps"], r["gapopen"]) == (2 * 80 - 15, 83, 80, 0, 3, 1)
    r = O.extend((x + y).encode(), (x + "T" + y[:20] + "C" + y[20:]).encode())   # two separate 1-base gaps
    assert (r["gaps"], r["gapopen"], r["qi"], r["sj"]) == (2, 2, 80, 82)
    assert O.extend(b"", b"ACGT")["score2"] == 0
    r = O.extend(b"ACGTNACGT", b"ACGTNACGT")                                  # N never matches
    assert (r["score2"], r["qi"], r["mismatch"]) == (12, 9, 1)


@pytest.mark.parametrize("kw", [
    dict(taxa=2, count=120, seed=1),
    dict(taxa=2, count=80, seed=2, mutation_rate=0.08),
    dict(taxa=2, count=80, seed=3, mutation_rate=0.03, indel_rate=0.004),
    dict(taxa=2, count=80, seed=4, mutation_rate=0.15, indel_rate=0.01),
])
def test_greedy_and_dp_agree_on_score_and_end_points(kw):
    from oracle import oracle as O
    _, raw = make_samples(**kw)
    s = [O.Sample(*r) for r in raw]
    try:
        O.set_gapped_mode(False)
        dp = O.search_pair(s[0], s[1]).hsps()
    finally:
        O.set_gapped_mode(True)
    gr = O.search_pair(s[0], s[1]).hsps()
    assert len(dp) == len(gr) and len(gr) > 0
    assert np.array_equal(dp[:, :8], gr[:, :8])          # transcripts, strand, ranges, score
    # differences (mismatches + gap characters) are equal too: score = t - 6d
    assert np.array_equal(dp[:, 8] + dp[:, 9], gr[:, 8] + gr[:, 9])
