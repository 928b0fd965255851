#!/usr/bin/env python
"""One stream, a few full steps of the HBM-resident all-vs-all: the command profiled under ncu for the launch
lists and --set full captures in profiles/ (bench.py itself drives six streams and several timed legs).

  python tools/profile_step.py [--config 1] [--steps 1] [--warmup 1] [--pairs N]"""
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
    ap.add_argument("--taxa", type=int, default=0)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=1)
    a = ap.parse_args()
    taxa, count, top = CONFIGS[a.config]
    taxa = a.taxa or taxa
    import torch
    from rna_clique_b200.engine import Engine, select_top_genes
    from rna_clique_b200.simulate import simulate
    _, ts = simulate(taxa, count, seed=487, device="cuda:0")
    eng = Engine(0, streams=1)
    inputs, labels = [], []
    for t in ts:
        offs = t.offsets.numpy()
        keep = select_top_genes(t.coverage, t.gene, top)
        inputs.append(dict(ascii=t.ascii, starts=offs[:-1][keep], lengths=np.diff(offs).astype(np.int32)[keep],
                           coverage=t.coverage[keep], gene=t.gene[keep], iso=t.iso[keep]))
        labels.append(f"od1/{t.name}_top.fasta")
    for _ in range(a.warmup + a.steps):
        res = eng.all_vs_all(inputs, labels, None)
    torch.cuda.synchronize()
    st = eng.stage_times()
    print({k: round(v["ms"] / (a.warmup + a.steps), 3) for k, v in st.items()}, "launches", eng.launches, "ideal", res["counts"]["ideal"])


if __name__ == "__main__":
    main()
