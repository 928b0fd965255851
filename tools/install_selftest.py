#!/usr/bin/env python
"""The reference's installation self-test as a workload (docs/testing_installation.md:19-22, tests/verify_install/
minimal_config.yaml:2-39): 8 simulated taxa x 1,000 transcripts of ~2,050 nt, rna-clique -n 1000 (8 samples, 28 pairs =
56 blastn searches upstream, graph, distance matrix) - "around one minute on a modern desktop with one thread" there.
Prints the wall time of rna_clique() from FASTA directories here.  python tools/install_selftest.py [--calls 3]"""
import argparse
import os
import shutil
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=3)
    a = ap.parse_args()
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import simulate, write_dataset
    _, ts = simulate(8, 1000, seed=487, device="cpu")
    tmp = Path(tempfile.mkdtemp(prefix="rcq_selftest_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None))
    try:
        dirs = write_dataset(tmp / "in", ts)
        for rep in range(1 + a.calls):
            for d in ("od1", "od2"):
                shutil.rmtree(tmp / d, ignore_errors=True)
                (tmp / d).mkdir()
            t0 = time.perf_counter()
            sim, _ = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "distance_matrix.h5", 1000)
            dt = time.perf_counter() - t0
            m = sim.get_dissimilarity_df()
            print(f"call {rep}{' (warm-up)' if rep == 0 else ''}: {dt:.3f} s, {m.shape[0]} samples, {sim.ideal_component_count} ideal components, "
                  f"largest distance {float(m.to_numpy().max()):.4f}")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
