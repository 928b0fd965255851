#!/usr/bin/env python
"""Is the CPU oracle's time linear in the number of genes per sample?  bench.py's CPU arm reports a RATE
(extension steps per second) measured on a bounded sample of the workload; the rate only stands for the full
workload if it does not depend on the sample size.  One sample pair of the bench workload's simulator at 1,500,
3,000 and 6,000 genes per sample: seconds, extension steps, rate.  CPU only.

  python tools/cpu_linearity.py [--taxa 10] [--sizes 1500 3000 6000]"""
import argparse
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=10)
    ap.add_argument("--sizes", type=int, nargs="+", default=[1500, 3000, 6000])
    ap.add_argument("--seed", type=int, default=487)
    a = ap.parse_args()
    from oracle import oracle as O
    from rna_clique_b200.simulate import simulate
    base = None
    for n in a.sizes:
        _, ts = simulate(a.taxa, n, seed=a.seed, device="cpu")
        raw = [(t.ascii.numpy(), t.offsets.numpy(), t.gene.astype(np.int32), t.iso.astype(np.int32)) for t in ts[:2]]
        t0 = time.perf_counter()
        s = [O.Sample(*r) for r in raw]
        t1 = time.perf_counter()
        r = O.search_pair(s[0], s[1])
        t2 = time.perf_counter()
        rate = r.counts["dp_cells"] / (t2 - t0)
        base = base or rate
        print(f"{n:6d} genes per sample: samples {t1 - t0:6.2f} s, search {t2 - t1:6.2f} s, {r.counts['dp_cells']:.3e} extension steps, "
              f"{rate / 1e6:7.2f} M steps/s ({rate / base:.2f} x the first)")


if __name__ == "__main__":
    main()
