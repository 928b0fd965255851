"""CPU: the oracle's table-assembly restatement against fixtures produced by the REFERENCE's own
HomologFinder.get_match_table (tests/golden/make_golden.py)."""
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
ROW = ["length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "score", "gaps", "nident",
       "qgene", "qiso", "sgene", "siso", "qlen", "minus", "reverse"]


def frame_to_rows(fr):
    n = len(fr["length"])
    out = np.zeros((n, 17), dtype=np.int32)
    for k, c in enumerate(ROW):
        if c == "score":
            out[:, k] = np.asarray(fr["bitscore"]).astype(np.int32)
        elif c == "minus":
            out[:, k] = [s == "minus" for s in fr["sstrand"]]
        elif c in ("qlen", "reverse"):
            out[:, k] = 0
        else:
            out[:, k] = fr[c]
    return out


@pytest.mark.parametrize("name", ["basic", "ties", "first", "sparse", "top2"])
def test_match_table_against_reference(name):
    from oracle import oracle as O
    rec = json.loads((GOLD / f"match_table_{name}.json").read_text())
    res = O.match_table_from_rows(frame_to_rows(rec["fwd"]), frame_to_rows(rec["bwd"]), rec["top_n"], rec["keep_all"])
    got = res.rows(2)
    exp = rec["expect"]
    want = np.zeros((len(exp["length"]), 17), dtype=np.int32)
    for k, c in enumerate(ROW):
        if c == "score":
            want[:, k] = np.asarray(exp["bitscore"]).astype(np.int32)
        elif c == "minus":
            want[:, k] = [s == "minus" for s in exp["sstrand"]]
        elif c == "qlen":
            want[:, k] = 0
        else:
            want[:, k] = np.asarray(exp[c]).astype(np.int32)
    assert got.shape == want.shape
    # same multiset of rows
    assert sorted(map(tuple, got.tolist())) == sorted(map(tuple, want.tolist()))
    # and the reference's row order
    assert np.array_equal(got, want)
