"""CPU, world size 2 (gloo): the table exchange (exchange.allgather_flat) and the pair schedule —
the host logic of the N > 1 path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rna_clique_b200.engine import schedule_pairs
        from rna_clique_b200.exchange import allgather_flat
        totals = [1000, 900, 800, 700, 50]
        pairs, owner = schedule_pairs(len(totals), totals, world)
        mine = [k for k in range(len(pairs)) if owner[k] == rank]
        # item k: a ragged int32 "table" whose content encodes k (item 3 is empty)
        make = lambda k: torch.arange(17 * ((k * 7) % 5 if k != 3 else 0), dtype=torch.int32) + 1000 * k
        got = allgather_flat(dist, mine, [make(k) for k in mine], len(pairs), torch.device("cpu"))
        ok = all(torch.equal(got[k], make(k)) for k in range(len(pairs)))
        # sharded build + broadcast (engine: every rank builds samples k % world == rank)
        from rna_clique_b200.exchange import share_built
        built_here = []
        payload = lambda k: (np.arange(100 * k + 3, dtype=np.int64) * (k + 1)).view(np.uint8).copy()

        def build(k):
            built_here.append(k)
            return ("own", k, payload(k))

        def export_to(obj, buf):
            buf[:len(obj[2])] = torch.from_numpy(obj[2])

        objs = share_built(dist, 5, build, lambda o: [o[1], 7, 0, 0, 0, 0, 0, len(o[2])], export_to,
                           lambda k, meta, buf: ("imported", int(meta[0]), buf.numpy().copy()), torch.device("cpu"))
        ok = ok and built_here == [k for k in range(5) if k % world == rank]
        ok = ok and all(o[1] == k and np.array_equal(o[2], payload(k)) for k, o in enumerate(objs))
        ok = ok and all((o[0] == "own") == (k % world == rank) for k, o in enumerate(objs))
        # pipelined form: one thread issues the rounds in order, another fetches items as they arrive
        import threading
        from rna_clique_b200.exchange import PipelinedShare
        sh = PipelinedShare(dist, 5, torch.device("cpu"))
        got2 = {}

        def consumer():
            for k in range(5):
                if k % world != rank:
                    got2[k] = sh.fetch(k, lambda k, meta, buf: ("imported", int(meta[0]), buf.numpy().copy()))

        th = threading.Thread(target=consumer)
        th.start()
        for j in range(sh.rounds):
            k = sh.own_item(j)
            obj = ("own", k, payload(k)) if k is not None else None
            sh.issue_round(j, obj, lambda o: [o[1], 7, 0, 0, 0, 0, 0, len(o[2])], export_to)
        th.join(timeout=60)
        sh.finish()
        ok = ok and sh.rounds == 3 and sorted(got2) == [k for k in range(5) if k % world != rank]
        ok = ok and all(o[1] == k and np.array_equal(o[2], payload(k)) for k, o in got2.items())
        q.put((rank, ok, mine))
    finally:
        dist.destroy_process_group()


def test_allgather_flat_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    owned = sorted(k for _, _, mine in res for k in mine)
    assert owned == list(range(10))           # every pair owned exactly once
    assert all(len(mine) >= 3 for _, _, mine in res)
