#!/usr/bin/env python
"""rna_clique() from FASTA directories with the per-phase trace (RCQ_TRACE_API=1): python tools/time_api.py [--config 1] [--calls 3]"""
import argparse
import os
import shutil
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
CONFIGS = {0: (3, 5000, 2000), 1: (10, 60000, 50000), 2: (30, 100000, 50000)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=1)
    ap.add_argument("--calls", type=int, default=3)
    ap.add_argument("--jobs", type=int, default=0)
    a = ap.parse_args()
    os.environ["RCQ_TRACE_API"] = "1"
    taxa, count, top = CONFIGS[a.config]
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import simulate, write_dataset
    _, ts = simulate(taxa, count, seed=487, device="cuda:0")
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    tmp = Path(tempfile.mkdtemp(prefix="rcq_api_", dir=base))
    try:
        t0 = time.perf_counter()
        dirs = write_dataset(tmp / "in", ts)
        print(f"dataset written in {time.perf_counter() - t0:.1f} s", file=sys.stderr)
        del ts
        jobs = a.jobs or max((os.cpu_count() or 2) - 1, 1)
        for rep in range(1 + a.calls):
            for d in ("od1", "od2"):
                shutil.rmtree(tmp / d, ignore_errors=True)
                (tmp / d).mkdir()
            t0 = time.perf_counter()
            sim, _ = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "distance_matrix.h5", top, jobs=jobs)
            print(f"call {rep}: {time.perf_counter() - t0:.3f} s  (jobs {jobs}, ideal components {sim.ideal_component_count})", file=sys.stderr)
            del sim
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
