"""Diagnostic: wall time of the reference-facing rna_clique() call on a simulated dataset (taxa count top)."""
import sys, time, os, tempfile
from pathlib import Path
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rna_clique_b200.simulate import simulate, write_dataset
from rna_clique_b200.rna_clique import rna_clique
taxa, count, top = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
tmp = Path(tempfile.mkdtemp())
_, ts = simulate(taxa, count, seed=487)
t0 = time.perf_counter(); dirs = write_dataset(tmp / "in", ts); print("write_dataset", round(time.perf_counter() - t0, 2), "s")
(tmp / "od1").mkdir(); (tmp / "od2").mkdir()
for rep in range(2):
    t0 = time.perf_counter()
    sim, pts = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "matrix.pkl", top)
    print("rna_clique()", rep, round(time.perf_counter() - t0, 2), "s; ideal components", sim.ideal_component_count)
