"""GPU: the N > 1 engine path (samples built by their owner and shipped in pipelined rounds, pairs sharded,
tables all-gathered, phase 2 on every rank) with two ranks sharing cuda:0 over gloo - NCCL refuses two
ranks on one device, and the test box has one GPU.  Both ranks must return the single-process result."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _inputs():
    from rna_clique_b200.simulate import simulate
    _, ts = simulate(5, 260, seed=77, mutation_rate=0.03, indel_rate=0.002)
    inputs, labels = [], []
    for t in ts:
        offs = t.offsets.numpy()
        inputs.append(dict(ascii=t.ascii.numpy(), starts=offs[:-1].copy(), lengths=np.diff(offs).astype(np.int32),
                           coverage=t.coverage, gene=t.gene, iso=t.iso))
        labels.append(f"od1/{t.name}_top.fasta")
    return inputs, labels


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rna_clique_b200.engine import Engine
        inputs, labels = _inputs()
        eng = Engine(0, streams=2)
        out = []
        for _ in range(2):      # twice: the second step reuses pools, side streams and cached buffers
            r = eng.all_vs_all(inputs, labels, top_genes=200, dist=dist)
            out.append((r["labels"], r["matrix"], r["nident_sum"].tolist(), r["denom_sum"].tolist(), r["counts"], r["n_local_pairs"]))
        eng.close()
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


def test_two_ranks_on_one_gpu_equal_the_single_process_run():
    import torch.multiprocessing as mp
    from rna_clique_b200.engine import Engine
    inputs, labels = _inputs()
    eng = Engine(0, streams=2)
    ref = eng.all_vs_all(inputs, labels, top_genes=200)
    eng.close()
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert ref["counts"]["ideal"] > 100
    for rank in range(world):
        for lab, m, ni, de, counts, n_local in res[rank]:
            assert lab == ref["labels"] and np.array_equal(m, ref["matrix"])
            assert ni == ref["nident_sum"].tolist() and de == ref["denom_sum"].tolist() and counts == ref["counts"]
    # pairs are drawn at run time: every pair is searched exactly once, by whichever rank was free
    for step in range(2):
        assert sum(res[rank][step][5] for rank in range(world)) == 10
