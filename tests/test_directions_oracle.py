"""The one-search-per-pair shortcut against two independent blastn-style directions (oracle only, CPU):
find_homologs.py:229-239 runs query sample 2 -> database sample 1 and the reverse; the product derives both
frames from one search.  tools/direction_asymmetry.py measures the difference at full size; here the quick
cases must keep every best-hit edge and move no distance by more than the north star tolerance (1e-3)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))


def test_shortcut_keeps_edges_and_distances(monkeypatch):
    import direction_asymmetry as D
    monkeypatch.setattr(sys, "argv", ["direction_asymmetry.py", "--quick"])
    for tot in D.main():
        assert tot["edges_diff"] == 0
        assert tot["dmax"] <= 1e-3
        assert tot["rows_diff"] <= max(0.3 * tot["rows"], 4)
