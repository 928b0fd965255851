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


def allgather_flat(dist, my_ids, my_chunks, n_total: int, device):
    """my_chunks[k]: 1-D int32 tensor owned by this rank for global item my_ids[k].
    Returns a list of n_total int32 tensors (every rank gets every item).
    Two collectives: sizes (all_gather of an int64 vector), then one padded payload all_gather."""
    world = dist.get_world_size()
    sizes = torch.zeros(n_total, dtype=torch.int64, device=device)
    for i, c in zip(my_ids, my_chunks):
        sizes[i] = c.numel()
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    # owner of item i = the rank that reported it (ids are disjoint across ranks)
    ids_by_rank = []
    for r in range(world):
        ids_by_rank.append(my_ids if r == dist.get_rank() else None)
    id_count = torch.tensor([len(my_ids)], dtype=torch.int64, device=device)
    counts = [torch.zeros_like(id_count) for _ in range(world)]
    dist.all_gather(counts, id_count)
    max_ids = int(max(int(c.item()) for c in counts))
    idv = torch.full((max(max_ids, 1),), -1, dtype=torch.int64, device=device)
    if my_ids:
        idv[:len(my_ids)] = torch.tensor(list(my_ids), dtype=torch.int64, device=device)
    all_ids = [torch.zeros_like(idv) for _ in range(world)]
    dist.all_gather(all_ids, idv)
    payload = torch.cat([c.reshape(-1) for c in my_chunks]) if my_chunks else torch.zeros(0, dtype=torch.int32, device=device)
    all_sizes = torch.stack(all_sizes).cpu().tolist()      # one device->host copy, then plain Python ints
    all_ids = [v.cpu().tolist() for v in all_ids]
    totals = [int(sum(all_sizes[r][i] for i in all_ids[r] if i >= 0)) for r in range(world)]
    pad = max(max(totals), 1)
    buf = torch.zeros(pad, dtype=torch.int32, device=device)
    buf[:payload.numel()] = payload
    gathered = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(gathered, buf)
    if device.type == "cuda":
        # the consumers (rcq_table_from_flat) run on the library's own streams, which are not ordered
        # after torch's: finish the collective on the host before handing out views of `gathered`
        torch.cuda.synchronize(device)
    out = [None] * n_total
    for r in range(world):
        off = 0
        for i in all_ids[r]:
            if i < 0:
                continue
            n = int(all_sizes[r][i])
            out[i] = gathered[r][off:off + n]
            off += n
    return out


def share_built(dist, n_items: int, build, meta_of, export_to, import_from, device, meta_len: int = 8):
    """Sharded build + broadcast.  Rank r builds the items k with k % world == r (`build(k)`), every rank
    ends up with all n_items: `meta_of(obj)` -> int64[meta_len] whose last entry is the flat byte size,
    `export_to(obj, uint8 tensor)` fills the flat buffer on the owner, `import_from(k, meta, uint8 tensor)`
    rebuilds the object on the other ranks.  Collectives: one all-reduce of the metadata rows (disjoint,
    so the sum is a gather) and one broadcast per item, all in flight together."""
    world, rank = dist.get_world_size(), dist.get_rank()
    mine = [k for k in range(n_items) if k % world == rank]
    built = {k: build(k) for k in mine}
    metas = torch.zeros((n_items, meta_len), dtype=torch.int64, device=device)
    for k, obj in built.items():
        metas[k] = torch.as_tensor(np.asarray(meta_of(obj), dtype=np.int64), device=device)
    dist.all_reduce(metas)
    metas = metas.cpu().numpy()
    bufs, works = [], []
    for k in range(n_items):
        buf = torch.empty(max(int(metas[k, meta_len - 1]), 1), dtype=torch.uint8, device=device)
        if k in built:
            export_to(built[k], buf)
        bufs.append(buf)
        works.append(dist.broadcast(buf, src=k % world, async_op=True))
    for w in works:
        w.wait()
    if device.type == "cuda":
        torch.cuda.synchronize(device)
    out = [built[k] if k in built else import_from(k, metas[k], bufs[k]) for k in range(n_items)]
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
