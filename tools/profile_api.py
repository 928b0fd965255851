import sys, os, cProfile, pstats, tempfile
from pathlib import Path
sys.path.insert(0, "/root/repo")
from rna_clique_b200.simulate import simulate, write_dataset
from rna_clique_b200.rna_clique import rna_clique
tmp = Path(tempfile.mkdtemp())
_, ts = simulate(10, 60000, seed=487)
dirs = write_dataset(tmp / "in", ts)
(tmp / "od1").mkdir(); (tmp / "od2").mkdir()
rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "matrix.pkl", 50000)
pr = cProfile.Profile(); pr.enable()
sim, pts = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "matrix.pkl", 50000)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
