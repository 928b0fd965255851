"""Diagnostic: wall time of rcq_sample_create for one configs[1]-sized sample, alone on the GPU."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rna_clique_b200 import _native
from rna_clique_b200.engine import select_top_genes
from rna_clique_b200.simulate import simulate

_, ts = simulate(1, 60000, seed=487, device="cuda:0")
t = ts[0]
offs = t.offsets.numpy()
keep = select_top_genes(t.coverage, t.gene, 50000)
starts, lengths = offs[:-1][keep].copy(), np.diff(offs).astype(np.int32)[keep]
ctx = _native.Context(0)
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s = _native.Sample(ctx, t.ascii, starts, t.gene[keep], t.iso[keep], lengths=lengths)
    torch.cuda.synchronize()
    print(f"build {rep}: {1e3 * (time.perf_counter() - t0):.3f} ms  n_index {s.n_index}")
    s.free() if hasattr(s, "free") else None
