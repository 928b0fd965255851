"""Relative orientation of the isoforms of one gene, with the device search instead of BLAST.

Mirrors `blast_pairwise_get_strands` (reference export_orthologs.py:285-329): all isoforms are searched
against themselves (blastn, evalue 1e-5), hits with qseqid > sseqid are kept, and for every (qseqid,
sseqid) pair the hit with the highest bitscore gives +1 (same strand) or -1 (reverse complement).  Here
the isoforms become one device sample searched against itself (kernel 3, `rcq_search_pair`); the gapped
HSP set of that search is reduced on the host.  Ties between a plus and a minus HSP go to the plus one
(BLAST's tabular order is not specified for that case)."""
from __future__ import annotations

from typing import Iterable, Iterator, Optional

import numpy as np

from . import _native


def _sequence_of(rec) -> str:
    seq = getattr(rec, "seq", rec)
    return str(seq)


def pairwise_get_strands(isoforms: Iterable, evalue: float = 1e-5, ctx: Optional[_native.Context] = None
                         ) -> Iterator[tuple]:
    """isoforms: iterable of (isoform id, record) with record.id and record.seq (a Bio.SeqRecord, or anything
    shaped like one).  Yields ((iso_q, id_q), (iso_s, id_s), +1 | -1) for id_q > id_s, as the reference does."""
    isoforms = list(isoforms)
    if len(isoforms) <= 1:
        return
    names = [rec.id for _, rec in isoforms]
    iso_of = {rec.id: iso for iso, rec in isoforms}
    seqs = [_sequence_of(rec).encode() for _, rec in isoforms]
    offsets = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offsets[1:])
    ascii_ = np.frombuffer(b"".join(seqs), dtype=np.uint8)
    own = ctx is None
    ctx = ctx or _native.Context(0)
    before = _native.debug_keep_hsps(ctx, True)
    try:
        sample = _native.Sample(ctx, ascii_, offsets, np.arange(len(seqs), dtype=np.int32), np.ones(len(seqs), dtype=np.int32))
        # the e-value cut is applied below per HSP; the table of the search itself is not used
        _native.search_pair(ctx, sample, sample, evalue=1.0).close()
        hsps = _native.debug_last_hsps(ctx)
        sample.close()
    finally:
        _native.debug_keep_hsps(ctx, before)
        if own:
            ctx.close()
    L = _native.lib()
    total, n = int(offsets[-1]), len(seqs)
    best = {}
    for tx1, tx2, minus, _q0, _q1, _s0, _s1, score2 in hsps[:, :8].tolist():
        q, s = names[tx2], names[tx1]          # forward search: query = second sample, database = first
        if not q > s:
            continue
        score = score2 // 2
        if L.rcq_evalue(int(score), int(offsets[tx2 + 1] - offsets[tx2]), total, n) > evalue:
            continue
        cur = best.get((q, s))
        if cur is None or score > cur[0] or (score == cur[0] and minus < cur[1]):
            best[(q, s)] = (score, minus)
    for (q, s), (_, minus) in sorted(best.items()):
        yield (iso_of[q], q), (iso_of[s], s), (-1 if minus else 1)
