"""Diagnostic: per-call host wall times of uploads and pair searches over a few steps (configs[1])."""
import sys, time, threading, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rna_clique_b200 import _native
from rna_clique_b200.engine import Engine, select_top_genes
from rna_clique_b200.simulate import simulate

streams = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
host_mode = len(sys.argv) > 3 and sys.argv[3] == "host"
_, ts = simulate(10, 60000, seed=487, device="cuda:0")
inputs, labels = [], []
for t in ts:
    offs = t.offsets.numpy()
    keep = select_top_genes(t.coverage, t.gene, 50000)
    if host_mode:
        inputs.append(dict(ascii=t.ascii.cpu().pin_memory(), starts=offs[:-1].copy(), lengths=np.diff(offs).astype(np.int32),
                           coverage=t.coverage, gene=t.gene, iso=t.iso))
        labels.append(t.name)
        continue
    inputs.append(dict(ascii=t.ascii, starts=offs[:-1][keep].copy(), lengths=np.diff(offs).astype(np.int32)[keep],
                       coverage=t.coverage[keep], gene=t.gene[keep], iso=t.iso[keep]))
    labels.append(t.name)
torch.cuda.synchronize()
eng = Engine(0, streams=streams)
log = []
orig_sp, orig_S = _native.search_pair, _native.Sample
def sp(*a, **k):
    t0 = time.perf_counter(); r = orig_sp(*a, **k); log.append(("search", threading.get_ident() % 1000, t0, time.perf_counter())); return r
class S(orig_S):
    def __init__(self, *a, **k):
        t0 = time.perf_counter(); super().__init__(*a, **k); log.append(("upload", threading.get_ident() % 1000, t0, time.perf_counter()))
_native.search_pair = sp
_native.Sample = S
for step in range(steps):
    log.clear()
    t0 = time.perf_counter()
    res = eng.all_vs_all(inputs, labels, 50000 if host_mode else None)
    t1 = time.perf_counter()
    d = sorted((e[3] - e[2]) * 1e3 for e in log if e[0] == "search")
    u = sorted((e[3] - e[2]) * 1e3 for e in log if e[0] == "upload")
    print(f"step {step}: {1e3*(t1-t0):.1f} ms  host {res['host_ms']}  search calls: med {d[len(d)//2]:.2f} max {d[-1]:.2f} sum {sum(d):.1f}; "
          f"upload: med {u[len(u)//2]:.2f} max {u[-1]:.2f}")
    slow = [e for e in log if (e[3] - e[2]) * 1e3 > 25]
    for e in slow[:12]:
        print("   slow", e[0], e[1], f"{(e[2]-t0)*1e3:.1f} -> {(e[3]-t0)*1e3:.1f}")
