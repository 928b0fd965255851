"""FASTA ingest for the hot path: one pass over the file bytes with numpy, no per-record Python
objects for the sequence data.  Replaces the Bio.SeqIO.parse passes of select_top_genes.py:110-127
on the host side; the letters go to the device as one flat uint8 buffer + (start, length)."""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path

import numpy as np


@dataclass
class Fasta:
    _headers: list         # header lines without '>' (id = first whitespace-separated token); None = decode lazily from raw
    ascii: np.ndarray      # uint8, all sequence letters concatenated, newlines removed
    offsets: np.ndarray    # int64 [n+1]
    raw: np.ndarray = None        # the file bytes and the byte span of every header line, kept so the
    hdr_start: np.ndarray = None  # default id pattern can be decoded natively (parse_fasta_ids) and the
    hdr_end: np.ndarray = None    # selected headers can be copied without becoming Python strings

    @property
    def headers(self):
        if self._headers is None:
            mv = memoryview(self.raw)
            self._headers = [str(mv[a + 1:b], "utf-8").rstrip("\r") for a, b in zip(self.hdr_start.tolist(), self.hdr_end.tolist())]
        return self._headers

    @property
    def ids(self):
        return [h.split(None, 1)[0] if h else "" for h in self.headers]

    def header_bytes(self, keep):
        """(uint8 header text of the records `keep` concatenated, int64 offsets [len(keep)+1]) without '>' or a
        trailing CR, gathered from the file bytes with numpy (no Python string per record)."""
        keep = np.asarray(keep, dtype=np.int64)
        if self.raw is None:
            hb = [self._headers[i].encode() for i in keep]
            off = np.zeros(len(hb) + 1, dtype=np.int64)
            np.cumsum([len(b) for b in hb], out=off[1:])
            return (np.frombuffer(b"".join(hb), dtype=np.uint8) if hb else np.zeros(0, np.uint8)), off
        a = self.hdr_start[keep] + 1
        b = self.hdr_end[keep].copy()
        has = b > a
        b[has] -= (self.raw[b[has] - 1] == 13)
        ln = np.maximum(b - a, 0)
        off = np.zeros(len(keep) + 1, dtype=np.int64)
        np.cumsum(ln, out=off[1:])
        idx = np.arange(int(off[-1]), dtype=np.int64) + np.repeat(a - off[:-1], ln)
        return self.raw[idx], off

    def __len__(self):
        return len(self.offsets) - 1


def read_fasta(path: Path) -> Fasta:
    """One native pass over the file bytes (rcq_fasta_scan / rcq_fasta_parse): letters with newlines, CR
    and spaces removed, record offsets, header spans.  Letters in front of the first header are kept in
    `ascii` (before offsets[0]) like the numpy statement below."""
    from . import _native
    raw = np.fromfile(str(path), dtype=np.uint8)
    if raw.size == 0:
        return Fasta([], np.zeros(0, np.uint8), np.zeros(1, np.int64))
    ascii_, offsets, hs, he = _native.fasta_parse(raw)
    return Fasta(None, ascii_, offsets, raw, hs, he)    # header strings are decoded only if somebody asks


def read_fasta_numpy(path: Path) -> Fasta:
    """The same parse written with numpy (tests compare the two)."""
    raw = np.fromfile(str(path), dtype=np.uint8)
    if raw.size == 0:
        return Fasta([], np.zeros(0, np.uint8), np.zeros(1, np.int64))
    nl = np.flatnonzero(raw == 10)
    line_start = np.concatenate([[0], nl + 1])
    line_start = line_start[line_start < raw.size]
    is_hdr = raw[line_start] == ord(">")
    hdr_start = line_start[is_hdr]
    # end of each header line
    ends = np.append(nl, raw.size)
    hdr_end = ends[np.searchsorted(ends, hdr_start)]
    headers = [raw[a + 1:b].tobytes().decode().rstrip("\r") for a, b in zip(hdr_start, hdr_end)]
    # keep = sequence letters: not newline / CR / space, and not inside a header line
    keep = (raw != 10) & (raw != 13) & (raw != 32)
    delta = np.zeros(raw.size + 1, dtype=np.int32)
    np.add.at(delta, hdr_start, 1)
    np.add.at(delta, np.minimum(hdr_end, raw.size), -1)
    keep &= np.cumsum(delta[:-1]) == 0
    csum = np.concatenate([[0], np.cumsum(keep, dtype=np.int64)])
    offsets = np.append(csum[hdr_start], csum[-1]).astype(np.int64)
    return Fasta(headers, np.ascontiguousarray(raw[keep]), offsets)


def write_fasta(path: Path, headers, ascii_: np.ndarray, starts, lengths, width: int = 60) -> None:
    """Biopython-style output (SeqIO.write(..., "fasta"): '>' + header, 60 letters per line), as
    select_top_genes_all.py:37-45 writes od1/<sample>_top.fasta.  Formatted natively in one buffer.
    headers: list of str, or the (bytes, offsets) pair of Fasta.header_bytes."""
    from . import _native
    _native.fasta_format(ascii_, starts, lengths, headers if isinstance(headers, tuple) else list(headers), width).tofile(str(path))


def write_fasta_python(path: Path, headers, ascii_: np.ndarray, starts, lengths, width: int = 60) -> None:
    """The same writer in Python (tests compare the two)."""
    with open(path, "wb") as f:
        for h, s, n in zip(headers, starts, lengths):
            seq = ascii_[int(s):int(s) + int(n)]
            f.write(b">" + h.encode() + b"\n")
            full = (len(seq) // width) * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = seq[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if len(seq) > full:
                f.write(seq[full:].tobytes() + b"\n")
