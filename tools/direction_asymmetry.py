#!/usr/bin/env python
"""How much does the one-search-per-pair shortcut change?  ("Not modelled".)

The reference runs two blastn searches per sample pair (find_homologs.py:229-239): query sample 2 against
database sample 1 and the reverse.  The product (and orc_search_pair) runs ONE search - sample 1's
transcripts in the query role of every extension - and derives both direction frames from that HSP set.
The oracle's orc_search_pair_directions runs the two searches with their own query roles.  This script
compares the two tables on BASELINE configs[0] (3 x 5,000 transcripts, top 2,000), on indel-heavy
simulations and on the multi-isoform / paralog families of the selection tests, and prints, per case:
rows, rows that differ, best-hit edges that differ, and the largest change of a pair's distance.

CPU only (oracle/ is test infrastructure).  Usage: python tools/direction_asymmetry.py [--quick]"""
import itertools
import sys
from fractions import Fraction
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def compare(name, raw):
    from oracle import oracle as O
    o = [O.Sample(*r) for r in raw]
    tot = dict(rows=0, rows_diff=0, edges=0, edges_diff=0, dmax=0.0)
    for i, j in itertools.combinations(range(len(o)), 2):
        a = O.search_pair(o[i], o[j]).rows(2)
        b = O.search_pair(o[i], o[j], independent_directions=True).rows(2)
        sa = {tuple(r) for r in a.tolist()}
        sb = {tuple(r) for r in b.tolist()}
        ea = {(r[10], r[12]) for r in a.tolist()}
        eb = {(r[10], r[12]) for r in b.tolist()}
        tot["rows"] += len(sb)
        tot["rows_diff"] += len(sa ^ sb)
        tot["edges"] += len(eb)
        tot["edges_diff"] += len(ea ^ eb)
        if len(a) and len(b):
            da = 1 - Fraction(int(a[:, 9].sum()), int((a[:, 0] - a[:, 8]).sum()))
            db = 1 - Fraction(int(b[:, 9].sum()), int((b[:, 0] - b[:, 8]).sum()))
            tot["dmax"] = max(tot["dmax"], abs(float(da - db)))
    print(f"{name:58s} rows {tot['rows']:7d}  differing {tot['rows_diff']:5d}  edges {tot['edges']:7d}  "
          f"differing {tot['edges_diff']:3d}  max |d distance| {tot['dmax']:.2e}")
    return tot


def main():
    quick = "--quick" in sys.argv
    from helpers import make_family_samples, make_samples
    from rna_clique_b200.select_top_genes import top_gene_mask_numpy
    from rna_clique_b200.simulate import simulate
    cases = []
    if not quick:
        _, ts = simulate(3, 5000, seed=487)
        raw = []
        for t in ts:
            keep = top_gene_mask_numpy(t.coverage, t.gene, 2000)
            a, offs = t.ascii.numpy(), t.offsets.numpy()
            ln = np.diff(offs)[keep]
            buf = np.concatenate([a[offs[k]:offs[k + 1]] for k in keep])
            raw.append((buf, np.concatenate([[0], np.cumsum(ln)]).astype(np.int64), t.gene[keep], t.iso[keep]))
        cases.append(("configs[0]: 3 x 5,000 transcripts, top 2,000", raw))
    n = 60 if quick else 400
    cases.append((f"indels 0.004, subs 0.03, 2 x {n}", make_samples(taxa=2, count=n, seed=3, mutation_rate=0.03, indel_rate=0.004)[1]))
    cases.append((f"indels 0.03, subs 0.05, 2 x {n}", make_samples(taxa=2, count=n, seed=5, mutation_rate=0.05, indel_rate=0.03)[1]))
    cases.append((f"indels 0.01, subs 0.15, 2 x {n}", make_samples(taxa=2, count=n, seed=4, mutation_rate=0.15, indel_rate=0.01)[1]))
    cases.append(("isoform / paralog families, 3 samples", make_family_samples(seed=1, genes=40 if quick else 150, taxa=3)))
    return [compare(n_, r) for n_, r in cases]


if __name__ == "__main__":
    main()
