"""Diagnostic: task timeline of one engine step (RCQ_TRACE_TASKS), per rank; run under torchrun for N > 1."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["RCQ_TRACE_TASKS"] = "1"
import torch
from rna_clique_b200.engine import Engine, select_top_genes
from rna_clique_b200.simulate import simulate

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
CONFIGS = {1: (10, 60000), 2: (30, 100000)}
taxa, count = CONFIGS[int(os.environ.get("RCQ_TRACE_CONFIG", "1"))]
summary = bool(os.environ.get("RCQ_TRACE_SUMMARY"))
_, ts = simulate(taxa, count, seed=487, device=f"cuda:{local}")
inputs, labels = [], []
for t in ts:
    offs = t.offsets.numpy()
    keep = select_top_genes(t.coverage, t.gene, 50000)
    inputs.append(dict(ascii=t.ascii, starts=offs[:-1][keep].copy(), lengths=np.diff(offs).astype(np.int32)[keep],
                       coverage=t.coverage[keep], gene=t.gene[keep], iso=t.iso[keep]))
    labels.append(t.name)
eng = Engine(local, streams=3)
for step in range(5):
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    res = eng.all_vs_all(inputs, labels, None, dist=dist)
for r in range(world):
    if dist is not None:
        dist.barrier()
    if r == rank and rank in (0, world - 1):
        print(f"--- rank {rank}: host {res['host_ms']}")
        if summary:
            kinds = {}
            for w, task, a, b in eng.task_trace:
                k = kinds.setdefault(task[0], dict(n=0, busy=0.0, first=1e9, last=0.0))
                k["n"] += 1; k["busy"] += b - a; k["first"] = min(k["first"], a); k["last"] = max(k["last"], b)
            for name, k in sorted(kinds.items()):
                print(f"  task {name}: {k['n']:4d} x, busy {k['busy']*1e3:8.1f} ms (all workers), first start {k['first']*1e3:7.1f}, last end {k['last']*1e3:7.1f}")
            for w0 in sorted({x[0] for x in eng.task_trace}):
                tt = sorted((a, b) for w, _, a, b in eng.task_trace if w == w0)
                busy = sum(b - a for a, b in tt)
                print(f"  worker {w0}: busy {busy*1e3:7.1f} ms of {tt[-1][1]*1e3:7.1f}")
            continue
        for w, task, a, b in sorted(eng.task_trace, key=lambda x: x[2]):
            print(f"  w{w} {task[0]}{task[1]:<3d} {a*1e3:7.2f} -> {b*1e3:7.2f}  ({(b-a)*1e3:5.2f} ms)")
        sys.stdout.flush()
if dist is not None:
    dist.destroy_process_group()
