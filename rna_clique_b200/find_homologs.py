"""One sample pair -> gene-matches table.  Mirrors the reference's find_homologs.py:176-302
(HomologFinder) with the two `blastn` subprocesses (find_homologs.py:124) and the pandas
reductions (:130, :215-302) replaced by ONE call into the CUDA library: rcq_search_pair.

Table contract (docs/formats.md:231-252): transcripts1 is the `ssample` side, so
`sgene/siso` name a transcript of transcripts1 and `qgene/qiso` one of transcripts2 in every row;
`reverse` marks rows that came from the search transcripts1 -> transcripts2, whose
qstart/qend/sstart/send/sstrand keep that search's own query/subject orientation."""
from __future__ import annotations

from pathlib import Path
from typing import Callable, Optional

import numpy as np
import pandas as pd

from . import _native
from .fasta import read_fasta
from .transcripts import default_parser, parse_fasta_ids

COLUMNS = ["pident", "length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "evalue", "bitscore",
           "gaps", "nident", "sstrand", "qgene", "qiso", "sgene", "siso", "reverse"]
KA_LAMBDA, KA_K = 1.28, 0.46   # blastn 1/-2 with linear gaps (csrc/common.cuh)
_INT_COLS = ["length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "gaps", "nident",
             "qgene", "qiso", "sgene", "siso"]


def _smallest_int(a: np.ndarray) -> np.ndarray:
    """pd.to_numeric(..., downcast="unsigned" / "signed"): the smallest integer dtype holding the values."""
    if a.size == 0 or a.dtype.kind not in "iu":
        return a
    lo, hi = int(a.min()), int(a.max())
    kinds = (np.uint8, np.uint16, np.uint32, np.uint64) if lo >= 0 else (np.int8, np.int16, np.int32, np.int64)
    for t in kinds:
        i = np.iinfo(t)
        if i.min <= lo and hi <= i.max:
            return a if a.dtype == t else a.astype(t)
    return a


def shrink_df(df: pd.DataFrame) -> pd.DataFrame:
    """Down-cast integer columns to the smallest width holding their values
    (reference shrink_df, find_homologs.py:58-80)."""
    for c in _INT_COLS:
        if c in df and len(df):
            df[c] = _smallest_int(df[c].to_numpy())
    return df


class SampleCache:
    """Device-resident samples keyed by FASTA path (the role of BlastDBCache, find_all_pairs.py:120-159)."""

    def __init__(self, ctx: Optional[_native.Context] = None, device: int = 0):
        self.ctx = ctx or _native.Context(device)
        self._samples = {}

    def get(self, path: Path, parser=default_parser) -> _native.Sample:
        path = Path(path)
        st = path.stat()
        key = (str(path.resolve()), st.st_mtime_ns, st.st_size, id(parser))
        s = self._samples.get(key)
        if s is None:
            fa = read_fasta(path)
            _, gene, iso = parse_fasta_ids(fa, parser)
            s = _native.Sample(self.ctx, fa.ascii, fa.offsets, gene, iso, name=str(path))
            self._samples[key] = s
        return s

    def preload(self, path: Path, parser, ascii_, offsets, gene, iso, lengths=None) -> None:
        """Builds the sample of a FASTA file this process has just written from the records it wrote, so
        that `get` does not read and parse the file again (same letters, offsets and ids by construction)."""
        path = Path(path)
        st = path.stat()
        key = (str(path.resolve()), st.st_mtime_ns, st.st_size, id(parser))
        if key not in self._samples:
            self._samples[key] = _native.Sample(self.ctx, ascii_, offsets, gene, iso, name=str(path), lengths=lengths)

    def close(self):
        for s in self._samples.values():
            s.close()
        self._samples.clear()


_default_cache: Optional[SampleCache] = None


def default_cache() -> SampleCache:
    global _default_cache
    if _default_cache is None:
        _default_cache = SampleCache()
    return _default_cache


def blast_printed_bitscore(b: np.ndarray) -> np.ndarray:
    """Bit scores as blastn's tabular output prints them (and simple_blast parses them back): above 99.9
    the value is TRUNCATED to an integer ("%3.0ld" of (long)bit_score), otherwise one decimal ("%4.1lf").
    Pinned by the real rows of the reference's docs/formats.md:260-273 (tests/golden/blast_rows_formats_md.json):
    nident 2309 / mismatch 19 -> 4194.86 bits -> printed 4194."""
    b = np.asarray(b, dtype=np.float64)
    return np.where(b > 99.9, np.floor(b), np.round(b, 1))


def blast_printed_evalue(e: np.ndarray) -> np.ndarray:
    """E-values as blastn's tabular output prints them: "0.0" below 1e-180, three significant digits
    ("%.2le") for small values (docs/formats.md:271 shows 8.75e-152), BLAST's fixed formats from 0.0009 up."""
    e = np.asarray(e, dtype=np.float64)
    out = e.copy()
    small = (e >= 1e-180) & (e < 0.0009)
    # round-trip through the decimal string: exactly what parsing BLAST's text yields
    if small.any():
        out[small] = [float(f"{x:.2e}") for x in e[small]]
    out[e < 1e-180] = 0.0
    for lo, hi, nd in ((0.0009, 0.1, 3), (0.1, 1.0, 2), (1.0, 10.0, 1), (10.0, np.inf, 0)):
        m = (e >= lo) & (e < hi)
        out[m] = np.round(e[m], nd)
    return out


def table_stats_numpy(cols: dict, n_db_bases_fwd: int, n_db_seqs_fwd: int, n_db_bases_bwd: int, n_db_seqs_bwd: int):
    """(pident, evalue, bitscore) as rcq_table_stats computes them, written with numpy (tests compare the two).
    Karlin-Altschul statistics; the length adjustment depends only on (query length,
    direction), so it is evaluated once per distinct value."""
    n = len(cols["length"])
    L = _native.lib()
    rev = np.asarray(cols["reverse"]).astype(bool)
    score, qlen = cols["score"], cols["qlen"]
    key = qlen.astype(np.int64) * 2 + rev
    uk, inv = np.unique(key, return_inverse=True)
    la = np.array([L.rcq_length_adjustment(int(k >> 1), *((n_db_bases_bwd, n_db_seqs_bwd) if k & 1 else
                                                         (n_db_bases_fwd, n_db_seqs_fwd))) for k in uk], dtype=np.float64)[inv] \
        if n else np.zeros(0)
    nb = np.where(rev, n_db_bases_bwd, n_db_bases_fwd).astype(np.float64)
    ns = np.where(rev, n_db_seqs_bwd, n_db_seqs_fwd).astype(np.float64)
    evalue = KA_K * np.maximum(qlen - la, 1.0) * np.maximum(nb - ns * la, 1.0) * np.exp(-KA_LAMBDA * score.astype(np.float64))
    evalue = blast_printed_evalue(evalue)
    bits = blast_printed_bitscore((KA_LAMBDA * score - np.log(KA_K)) / np.log(2.0))
    return np.round(100.0 * cols["nident"] / np.maximum(cols["length"], 1), 3), evalue, bits


def table_to_frame(cols: dict, n_db_bases_fwd: int, n_db_seqs_fwd: int, n_db_bases_bwd: int,
                   n_db_seqs_bwd: int) -> pd.DataFrame:
    """Device columns -> the reference's DataFrame.  pident / evalue / bitscore are derived here in
    fp64 from the integer columns (blastn tabular semantics: pident = 100 nident / length;
    bitscore printed as an integer above 99.9 bits)."""
    n = len(cols["length"])
    pident, evalue, bits = np.empty(n), np.empty(n), np.empty(n)
    c32 = {c: np.ascontiguousarray(cols[c], dtype=np.int32) for c in ("score", "qlen", "nident", "length")}
    rev8 = np.ascontiguousarray(cols["reverse"], dtype=np.uint8)
    _native.check(_native.lib().rcq_table_stats(n, c32["score"].ctypes.data, c32["qlen"].ctypes.data, rev8.ctypes.data,
                                                c32["nident"].ctypes.data, c32["length"].ctypes.data, int(n_db_bases_fwd),
                                                int(n_db_seqs_fwd), int(n_db_bases_bwd), int(n_db_seqs_bwd), float(np.log(KA_K)),
                                                float(np.log(2.0)), pident.ctypes.data, evalue.ctypes.data, bits.ctypes.data))
    rev = rev8.astype(bool)
    ic = {c: _smallest_int(np.asarray(cols[c])) for c in _INT_COLS}      # shrink_df, before the frame is built
    strands = pd.Series(["plus", "minus"])                                  # the frame's own string dtype
    df = pd.DataFrame({
        "pident": pident,
        "length": ic["length"], "mismatch": ic["mismatch"], "gapopen": ic["gapopen"],
        "qstart": ic["qstart"], "qend": ic["qend"], "sstart": ic["sstart"], "send": ic["send"],
        "evalue": evalue, "bitscore": bits, "gaps": ic["gaps"], "nident": ic["nident"],
        "sstrand": strands.take(cols["minus"].astype(np.int64)).reset_index(drop=True),
        "qgene": ic["qgene"], "qiso": ic["qiso"], "sgene": ic["sgene"], "siso": ic["siso"],
        "reverse": rev,
    }, columns=COLUMNS, copy=False)      # the arrays are this call's own: no defensive copy, no consolidation (16 ms of 28)
    return df


class HomologFinder:
    """HomologFinder(parse_transcript_id, top_n, evalue, keep_all, debug=False, **blast_kwargs)
    (find_homologs.py:176-213).  blast_kwargs are accepted for signature compatibility; `db_cache`
    may carry a SampleCache (what find_all_pairs passes), everything else is ignored."""

    def __init__(self, parse_transcript_id: Callable = default_parser, top_n: int = 1, evalue: float = 1e-99,
                 keep_all: bool = True, debug: bool = False, **blast_kwargs):
        self.parse, self.top_n, self.evalue, self.keep_all, self.debug = parse_transcript_id, top_n, evalue, keep_all, debug
        cache = blast_kwargs.get("db_cache")
        self.cache = cache if isinstance(cache, SampleCache) else default_cache()
        self.last_table: Optional[_native.Table] = None

    def get_device_table(self, transcripts1: Path, transcripts2: Path) -> _native.Table:
        s1, s2 = self.cache.get(transcripts1, self.parse), self.cache.get(transcripts2, self.parse)
        t = _native.search_pair(self.cache.ctx, s1, s2, self.evalue, self.top_n, self.keep_all)
        t.db = (s1.total, s1.n_seq, s2.total, s2.n_seq)
        return t

    def get_match_table(self, transcripts1: Path, transcripts2: Path) -> pd.DataFrame:
        t = self.get_device_table(transcripts1, transcripts2)
        self.last_table = t
        # forward search: query transcripts2 vs database transcripts1; backward: the other way round
        return table_to_frame(t.download(), t.db[0], t.db[1], t.db[2], t.db[3])
