"""CPU: the phase-2 oracle against fixtures produced by the REFERENCE's own build_graph +
SampleSimilarity (tests/golden/make_golden.py)."""
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
CASES = sorted(p.stem[len("phase2_"):] for p in GOLD.glob("phase2_*.json"))


def load_case(name):
    rec = json.loads((GOLD / f"phase2_{name}.json").read_text())
    names = rec["names"]
    idx = {s: i for i, s in enumerate(names)}
    tables = []
    for t in rec["tables"]:
        tables.append(dict(s=idx[t["ssample"]], q=idx[t["qsample"]],
                           **{c: np.asarray(t[c], dtype=np.int32) for c in ("sgene", "qgene", "nident", "length", "gaps")}))
    return rec, names, tables


@pytest.mark.parametrize("name", CASES)
def test_phase2_oracle_against_reference(name):
    from oracle import phase2_ref
    rec, names, tables = load_case(name)
    exp = rec["expect"]
    res = phase2_ref.phase2(len(names), tables, rec.get("sample_count"))
    assert res["sample_count"] == exp["sample_count"]
    assert res["n_nodes"] == exp["n_nodes"] and res["n_edges"] == exp["n_edges"]
    got_valid = sorted([names[s], g] for s, g in res["valid"])
    assert got_valid == exp["valid"]
    if exp["error"]:
        with pytest.raises(ZeroDivisionError):
            phase2_ref.distance_matrix(names, tables, res["sums"])
        return
    fr = {frozenset(k): (n, d) for k, n, d in exp["fractions"]}
    from fractions import Fraction
    for t, (ni, de) in zip(tables, res["sums"]):
        n, d = fr[frozenset((names[t["s"]], names[t["q"]]))]
        assert Fraction(ni, de) == Fraction(n, d)
    labels, m = phase2_ref.distance_matrix(names, tables, res["sums"])
    assert labels == exp["samples"]
    assert np.array_equal(m, np.asarray(exp["matrix"]))   # bit-exact fp64
