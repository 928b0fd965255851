#!/usr/bin/env python
"""Times the gapped stage alone (one stream) on a few sample pairs of the bench workload: ms per pair and GCUPS.
Used to compare builds of the library (RCQ_LIB=...).  python tools/solo_gapped.py [--config 1] [--pairs 6]"""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
CONFIGS = {0: (3, 5000, 2000), 1: (10, 60000, 50000), 2: (30, 100000, 50000)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=1)
    ap.add_argument("--pairs", type=int, default=6)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--variants", default="", help="comma-separated RCQ_GAPPED_VARIANT values to time in turn (8, 16)")
    a = ap.parse_args()
    taxa, count, top = CONFIGS[a.config]
    import itertools
    from rna_clique_b200 import _native
    from rna_clique_b200.engine import select_top_genes
    from rna_clique_b200.simulate import simulate
    _, ts = simulate(taxa, count, seed=487, device="cuda:0")
    ctx = _native.Context(0)
    allp = list(itertools.combinations(range(taxa), 2))
    pairs = [allp[k] for k in np.linspace(0, len(allp) - 1, num=min(a.pairs, len(allp))).astype(int)]
    need = sorted({k for p in pairs for k in p})
    ss = {}
    for k in need:
        t = ts[k]
        offs = t.offsets.numpy()
        keep = select_top_genes(t.coverage, t.gene, top)
        ss[k] = _native.Sample(ctx, t.ascii, offs[:-1][keep], t.gene[keep], t.iso[keep], lengths=np.diff(offs).astype(np.int32)[keep])
    import os
    for v in (a.variants.split(",") if a.variants else [None]):
        if v is not None:
            os.environ["RCQ_GAPPED_VARIANT"] = v
        _native.search_pair(ctx, ss[pairs[0][0]], ss[pairs[0][1]]).close()
        ctx.stage_times(reset=True)
        cells = retries = 0
        for _ in range(a.reps):
            for i, j in pairs:
                t = _native.search_pair(ctx, ss[i], ss[j])
                cells += t.stats["dp_cells"]; retries += t.stats["band_retries"]
                t.close()
        st = ctx.stage_times(reset=True)
        n = a.reps * len(pairs)
        g = st["gapped_dp"]["ms"] / n
        print(f"lib {_native.LIB_PATH.name} variant {v}: gapped {g:.3f} ms per pair, {cells / n / (g * 1e-3) / 1e9:.1f} GCUPS, {retries / n:.0f} groups given up per pair; "
              + ", ".join(f"{k} {v2['ms'] / n:.3f}" for k, v2 in st.items() if k not in ("gapped_dp", "pack_index", "phase2")))

if __name__ == "__main__":
    main()
