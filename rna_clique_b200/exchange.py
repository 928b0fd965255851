"""Multi-GPU exchange: the ONLY collective of the path.  Every rank searched its
own shard of the sample pairs; the flat int32 hit tables are all-gathered (torch.distributed:
NCCL over NVLink on the GPU box, gloo in the CPU tests) so that every rank holds the global edge
list and runs the component filter redundantly — there is no second collective."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native

NCOLS = 17


def _is_nccl(dist) -> bool:
    """The single-tensor all-gather is NCCL's; gloo (CPU tests, and the two-ranks-on-one-GPU test) takes lists."""
    try:
        return str(dist.get_backend()).lower() == "nccl"
    except Exception:  # noqa: BLE001
        return False


def allgather_flat(dist, my_ids, my_chunks, n_total: int, device):
    """my_chunks[k]: 1-D int32 tensor owned by this rank for global item my_ids[k].
    Returns a list of n_total int32 tensors (every rank gets every item).
    Two collectives: one all-reduce of (size, owner + 1) per item - ids are disjoint across ranks, so the
    sum is a gather - and one padded payload all_gather."""
    world, rank = dist.get_world_size(), dist.get_rank()
    info = np.zeros((2, n_total), dtype=np.int64)
    for i, c in zip(my_ids, my_chunks):
        info[0, i] = c.numel()
        info[1, i] = rank + 1
    info_t = torch.as_tensor(info, device=device)
    dist.all_reduce(info_t)
    info = info_t.cpu().numpy()                              # the one device->host copy of the exchange
    sizes, owner = info[0], info[1] - 1
    if (owner < 0).any() or (owner >= world).any():
        raise RuntimeError("allgather_flat: an item is owned by no rank or by several")
    ids_by_rank = [np.flatnonzero(owner == r) for r in range(world)]        # ascending ids per rank
    totals = [int(sizes[ids].sum()) for ids in ids_by_rank]
    pad = max(max(totals), 1)
    buf = torch.empty(pad, dtype=torch.int32, device=device)
    off = 0
    for i, c in sorted(zip(my_ids, my_chunks), key=lambda p: p[0]):
        buf[off:off + c.numel()] = c.reshape(-1)
        off += c.numel()
    gathered = torch.empty((world, pad), dtype=torch.int32, device=device)
    if _is_nccl(dist):
        dist.all_gather_into_tensor(gathered, buf)
    else:
        dist.all_gather(list(gathered.unbind(0)), buf)
    if device.type == "cuda":
        # the consumers (rcq_table_from_flat) run on the library's own streams, which are not ordered
        # after torch's: finish the collective on the host before handing out views of `gathered`
        torch.cuda.synchronize(device)
    out = [None] * n_total
    for r in range(world):
        off = 0
        for i in ids_by_rank[r]:
            n = int(sizes[i])
            out[int(i)] = gathered[r, off:off + n]
            off += n
    return out


class PipelinedShare:
    """Sharded build + broadcast, pipelined by rounds.  Item k belongs to rank k % world; round j covers the
    items j*world .. (j+1)*world-1.  ONE thread per rank calls `issue_round(j, obj, ...)` for j = 0, 1, ...
    with the item it built for that round (the collectives of all ranks therefore line up: per round one
    all-reduce of the metadata rows - disjoint, so the sum is a gather - and one asynchronous broadcast per
    item).  Any thread may then `fetch(k, import_from)`, which blocks until item k has arrived and rebuilds
    the object, so consumers start on the first items while later rounds are still being built and sent.
    meta_of(obj) -> int64[meta_len] whose last entry is the flat byte size; export_to(obj, uint8 tensor)
    fills the flat buffer on the owner; import_from(k, meta, uint8 tensor) rebuilds it elsewhere."""

    def __init__(self, dist, n_items: int, device, meta_len: int = 8, meta_group=None):
        import threading
        self.dist, self.n, self.device, self.meta_len = dist, n_items, device, meta_len
        # the metadata all-reduce may run on its own process group (own NCCL stream): on the default one it
        # would queue behind the payload of the previous round and hold back the next round's sends
        self.meta_group = meta_group
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.rounds = (n_items + self.world - 1) // self.world
        self.metas = [None] * n_items
        self.bufs = [None] * n_items
        self.works = [None] * n_items
        self.issued = [threading.Event() for _ in range(n_items)]
        self._side = threading.local()
        self._keep = []

    def own_item(self, j: int):
        k = j * self.world + self.rank
        return k if k < self.n else None

    def is_issued(self, k: int) -> bool:
        return self.issued[k].is_set()

    def arrived(self, k: int) -> bool:
        """Item k's collective has finished on the device (non-blocking): `fetch(k)` will not wait."""
        return self.issued[k].is_set() and bool(self.works[k].is_completed())

    def issue_round(self, j: int, obj, meta_of, export_to) -> None:
        lo, hi = j * self.world, min(self.n, (j + 1) * self.world)
        own = self.own_item(j)
        metas = torch.zeros((hi - lo, self.meta_len), dtype=torch.int64, device=self.device)
        if own is not None:
            metas[own - lo] = torch.as_tensor(np.asarray(meta_of(obj), dtype=np.int64), device=self.device)
        self.dist.all_reduce(metas, group=self.meta_group)
        metas = metas.cpu().numpy()
        sizes = [max(int(metas[k - lo, self.meta_len - 1]), 1) for k in range(lo, hi)]
        if hi - lo == self.world and self.world > 1 and _is_nccl(self.dist):
            # a full round: every rank contributes one item, so the round is ONE all-gather (all NVLink ports
            # busy in both directions) instead of `world` broadcasts queued behind each other
            pad = max(sizes)
            out = torch.empty((self.world, pad), dtype=torch.uint8, device=self.device)
            inp = torch.empty(pad, dtype=torch.uint8, device=self.device)
            export_to(obj, inp)              # synchronous on the library's stream: the data is there before the send
            work = self.dist.all_gather_into_tensor(out, inp, async_op=True)
            for k in range(lo, hi):
                self.metas[k], self.bufs[k], self.works[k] = metas[k - lo], out[k - lo, :sizes[k - lo]], work
                self.issued[k].set()
            self._keep.append(inp)
            return
        for k in range(lo, hi):
            buf = torch.empty(sizes[k - lo], dtype=torch.uint8, device=self.device)
            if k == own:
                export_to(obj, buf)
            self.metas[k], self.bufs[k] = metas[k - lo], buf
            self.works[k] = self.dist.broadcast(buf, src=k % self.world, async_op=True)
            self.issued[k].set()

    def fetch(self, k: int, import_from):
        self.issued[k].wait()
        w = self.works[k]
        if self.device.type == "cuda":
            # the consumer runs on the library's own streams, which are not ordered after torch's: make a
            # side stream wait for the collective and finish that stream on the host
            side = getattr(self._side, "stream", None)
            if side is None:
                side = self._side.stream = torch.cuda.Stream(self.device)
            with torch.cuda.stream(side):
                w.wait()
            side.synchronize()
        else:
            w.wait()
        obj = import_from(k, self.metas[k], self.bufs[k])   # synchronous copy out of the flat buffer
        self.bufs[k] = None
        return obj

    def finish(self) -> None:
        """Owners keep their send buffers until every broadcast has completed."""
        for k, w in enumerate(self.works):
            if w is not None and self.bufs[k] is not None:
                w.wait()
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self.bufs = [None] * self.n
        self._keep = []


def share_built(dist, n_items: int, build, meta_of, export_to, import_from, device, meta_len: int = 8):
    """Sharded build + broadcast, eager form: rank r builds the items k with k % world == r (`build(k)`),
    every rank returns all n_items (see PipelinedShare for the callbacks)."""
    sh = PipelinedShare(dist, n_items, device, meta_len)
    out = [None] * n_items
    for j in range(sh.rounds):
        k = sh.own_item(j)
        obj = build(k) if k is not None else None
        if k is not None:
            out[k] = obj
        sh.issue_round(j, obj, meta_of, export_to)
    for k in range(n_items):
        if out[k] is None:
            out[k] = sh.fetch(k, import_from)
    sh.finish()
    return out


def allgather_tables(ctx, dist, pairs, owner, mine, my_tables):
    """All ranks end up with one device table per pair, in `pairs` order."""
    device = torch.device("cuda", ctx.device)
    L = _native.lib()
    chunks = []
    for t in my_tables:
        n = int(L.rcq_table_flat_words(t.h))
        buf = torch.empty(max(n, 1), dtype=torch.int32, device=device)
        _native.check(L.rcq_table_flatten(ctx.h, t.h, buf.data_ptr()))
        chunks.append(buf[:n])
    flat = allgather_flat(dist, mine, chunks, len(pairs), device)
    mine_set = {k: t for k, t in zip(mine, my_tables)}
    tables = []
    for k in range(len(pairs)):
        if k in mine_set:
            tables.append(mine_set[k])
            continue
        words = flat[k].contiguous()
        h = C.c_void_p()
        _native.check(L.rcq_table_from_flat(ctx.h, words.data_ptr(), words.numel() // NCOLS, C.byref(h)))
        tables.append(_native.Table(ctx, h, None))
    return tables


def phase2_exchanged(ctx, dist, n_samples: int, stride: int, pairs, owner, mine, my_tables, sample_count=None,
                     want_valid: bool = False):
    """Phase 2 over the tables of all ranks without rebuilding them: every rank packs the five columns phase 2 reads
    (sgene, qgene, nident, length, gaps: 20 bytes per row instead of the 68 of a whole table) of its own tables into
    one block, the blocks are all-gathered (one collective for the row counts, one for the payload), and
    rcq_phase2_packed reduces the gathered buffer in place on every rank.  The library's stream waits for the
    collective through an event: no device-wide synchronisation."""
    import os
    import time
    world, rank = dist.get_world_size(), dist.get_rank()
    device = torch.device("cuda", ctx.device) if torch.cuda.is_available() and _is_nccl(dist) else torch.device("cpu")
    on_gpu = device.type == "cuda"
    trace = os.environ.get("RCQ_TRACE_EXCHANGE") and rank == 0
    marks = [("start", time.perf_counter())]

    def mark(name):
        if trace:
            if on_gpu:
                torch.cuda.synchronize(device)
            marks.append((name, time.perf_counter()))
    info = torch.zeros((2, len(pairs)), dtype=torch.int64, device=device)   # rows of every table, rank + 1 of its holder
    mine_sorted = sorted(range(len(mine)), key=lambda x: mine[x])          # local tables in global pair order
    local = [my_tables[x] for x in mine_sorted]
    if mine:
        ids = torch.as_tensor([mine[x] for x in mine_sorted], device=device)
        info[0, ids] = torch.as_tensor([len(t) for t in local], dtype=torch.int64, device=device)
        info[1, ids] = rank + 1
    dist.all_reduce(info)                                                  # disjoint ids: the sum is a gather
    info = info.cpu().numpy()
    rows, holder = info[0], info[1] - 1
    mark("row counts")
    if (holder < 0).any() or (holder >= world).any():
        raise RuntimeError("phase2_exchanged: a pair was searched by no rank or by several")
    if owner is not None and not np.array_equal(np.asarray(owner), holder):
        raise RuntimeError("phase2_exchanged: tables are not where the schedule put them")
    owner = holder
    n_by_rank = [int(rows[owner == r].sum()) for r in range(world)]
    pad = max(max(n_by_rank), 1)
    cuda_dev = torch.device("cuda", ctx.device)
    buf = torch.zeros(5 * pad, dtype=torch.int32, device=cuda_dev)
    _native.phase2_pack(ctx, local, buf.data_ptr())                        # synchronous on the library's stream
    mark("pack")
    gathered = torch.empty((world, 5 * pad), dtype=torch.int32, device=device)
    send = buf if buf.device == device else buf.to(device)
    if _is_nccl(dist):
        dist.all_gather_into_tensor(gathered, send)
    else:
        dist.all_gather(list(gathered.unbind(0)), send)
    mark("all-gather")
    if gathered.is_cuda:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(gathered.device))
        torch.cuda.ExternalStream(ctx.stream, device=gathered.device).wait_event(ev)
        dev = gathered
    else:
        dev = gathered.to(cuda_dev)                                        # gloo test path: the payload travelled over the host
    row_off = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    src_base = np.zeros(len(pairs), dtype=np.int64)
    src_stride = np.zeros(len(pairs), dtype=np.int64)
    at = [0] * world
    for k in range(len(pairs)):
        r = int(owner[k])
        src_base[k] = r * 5 * pad + at[r]
        src_stride[k] = n_by_rank[r]
        at[r] += int(rows[k])
    res = _native.phase2_packed(ctx, n_samples, stride, row_off, [p[0] for p in pairs], [p[1] for p in pairs], src_base, src_stride,
                                dev.data_ptr(), sample_count, want_valid)
    mark("phase 2")
    if trace:
        import sys
        print("[phase2_exchanged] " + ", ".join(f"{b[0]} {1e3 * (b[1] - a[1]):.2f} ms" for a, b in zip(marks, marks[1:])) +
              f"; {int(rows.sum())} rows, pad {pad}", file=sys.stderr)
    res["_keep"] = (dev, gathered, buf)
    return res
