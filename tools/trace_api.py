#!/usr/bin/env python
"""Task timeline of rna_clique() (RCQ_TRACE_TASKS): when samples become resident and when searches run.
python tools/trace_api.py [--config 2]"""
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
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--calls", type=int, default=2)
    a = ap.parse_args()
    os.environ["RCQ_TRACE_API"] = "1"
    os.environ["RCQ_TRACE_TASKS"] = "1"
    taxa, count, top = CONFIGS[a.config]
    from rna_clique_b200 import filtering_step as fs
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import simulate, write_dataset
    _, ts = simulate(taxa, count, seed=487, device="cuda:0")
    tmp = Path(tempfile.mkdtemp(prefix="rcq_api_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None))
    try:
        dirs = write_dataset(tmp / "in", ts)
        del ts
        jobs = max((os.cpu_count() or 2) - 1, 1)
        for rep in range(1 + a.calls):
            for d in ("od1", "od2"):
                shutil.rmtree(tmp / d, ignore_errors=True)
                (tmp / d).mkdir()
            t0 = time.perf_counter()
            sim, _ = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "distance_matrix.h5", top, jobs=jobs)
            print(f"call {rep}: {time.perf_counter() - t0:.3f} s", file=sys.stderr)
            tr = fs.shared_engine(jobs).task_trace
            ups = sorted((b, a_, x[1]) for w, x, a_, b in tr if x[0] == "u")
            ss = sorted((a_, b) for w, x, a_, b in tr if x[0] == "s")
            print("  uploads done at (s): " + " ".join(f"{b:.2f}" for b, _, _ in ups), file=sys.stderr)
            print(f"  upload task time: mean {sum(b - a_ for b, a_, _ in ups) / len(ups) * 1e3:.1f} ms", file=sys.stderr)
            n = len(ss)
            print(f"  searches: first start {ss[0][0]:.2f}, 25% {ss[n // 4][0]:.2f}, 50% {ss[n // 2][0]:.2f}, 75% {ss[3 * n // 4][0]:.2f}, "
                  f"last end {max(b for _, b in ss):.2f}; mean task {sum(b - a_ for a_, b in ss) / n * 1e3:.1f} ms", file=sys.stderr)
            del sim
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
