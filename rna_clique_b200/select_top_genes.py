"""Top-n genes by k-mer coverage.  Mirrors the reference's select_top_genes.py:36-137 and
select_top_genes_all.py:12-46: a gene's coverage is the maximum over its isoforms (starting from
0.0), the top n genes are heapq.nlargest over (coverage, gene) — so equal coverage resolves to the
LARGER gene id — and every isoform of a selected gene is emitted in input-file order."""
from __future__ import annotations

from pathlib import Path
from typing import Callable, Iterator

import numpy as np

from .fasta import Fasta, read_fasta, write_fasta
from .transcripts import default_parser, parse_fasta_ids


def top_gene_mask(coverage: np.ndarray, gene: np.ndarray, top: int) -> np.ndarray:
    """Indices (file order) of every transcript whose gene is among the `top` genes
    (select_top_genes.py:108-127).  Host code of the native library (rcq_top_gene_mask): one pass for
    the per-gene maxima and an nth_element, ~10x faster than the numpy statement below."""
    from . import _native
    return _native.top_gene_mask(coverage, gene, top)


def top_gene_mask_numpy(coverage: np.ndarray, gene: np.ndarray, top: int) -> np.ndarray:
    """The same selection written with numpy (tests compare the two)."""
    coverage = np.asarray(coverage, dtype=np.float64)
    gene = np.asarray(gene)
    ug, inv = np.unique(gene, return_inverse=True)
    best = np.zeros(len(ug), dtype=np.float64)       # defaultdict(float): starts at 0.0
    np.maximum.at(best, inv, coverage)
    if top >= len(ug):
        keep_gene = np.ones(len(ug), dtype=bool)
    elif top <= 0:
        keep_gene = np.zeros(len(ug), dtype=bool)
    else:
        order = np.lexsort((ug, best))               # ascending by coverage, then gene id
        keep_gene = np.zeros(len(ug), dtype=bool)
        keep_gene[order[len(ug) - top:]] = True
    return np.flatnonzero(keep_gene[inv])


class TopGeneSelector:
    """seqs_fun() -> Fasta (or a path-bound reader); top = number of genes; parser = id parser."""

    def __init__(self, seqs_fun: Callable[[], Fasta], top: int, parser=default_parser):
        self.seqs_fun, self.top, self.parser = seqs_fun, top, parser
        self._fa = None

    def _load(self):
        if self._fa is None:
            self._fa = self.seqs_fun()
            self._cov, self._gene, self._iso = parse_fasta_ids(self._fa, self.parser)
            self._keep = top_gene_mask(self._cov, self._gene, self.top)
        return self._fa

    def get_top_genes(self) -> Iterator[int]:
        self._load()
        return iter(sorted({int(g) for g in self._gene[self._keep]}))

    def get_top_gene_seqs(self):
        """(Fasta, kept transcript indices in file order, coverage, gene, isoform of the kept ones)."""
        fa = self._load()
        k = self._keep
        return fa, k, self._cov[k], self._gene[k], self._iso[k]

    @classmethod
    def from_path(cls, path: Path, *args, **kwargs):
        return cls(lambda: read_fasta(Path(path)), *args, **kwargs)


def _parse_and_select(out_dir: Path, transcripts: str, x: Path, *args):
    """Read, parse and select (thread-safe: no device work): returns (out path, x.stem, records, write), records =
    (parser, letters, starts, gene, iso, lengths) of the selected transcripts and write() = the callable that
    formats and writes out_dir/<x.stem>_top.fasta.  The device only needs `records`, so a pipeline can upload the
    sample while (or before) the file is written."""
    x = Path(x)
    out = Path(out_dir) / (x.stem + "_top.fasta")
    sel = TopGeneSelector.from_path(x / transcripts, *args)
    fa, keep, _, gene, iso = sel.get_top_gene_seqs()
    starts, lengths = fa.offsets[:-1][keep], np.diff(fa.offsets)[keep]

    def write():
        write_fasta(out, fa.header_bytes(keep), fa.ascii, starts, lengths)

    return out, x.stem, (sel.parser, fa.ascii, starts, np.asarray(gene, dtype=np.int32), np.asarray(iso, dtype=np.int32),
                         lengths.astype(np.int32)), write


def _parse_and_select_lean(out_dir: Path, transcripts: str, x: Path, top: int, parser=default_parser, alloc=None):
    """_parse_and_select for the default id pattern without reading what is thrown away: the file is mapped, not
    copied; ids come from the header lines; only the letters of the selected records are extracted, into a buffer
    from alloc(nbytes) (pinned memory in the pipeline, so that the upload is one DMA transfer).  Falls back to
    _parse_and_select for any other parser and for empty files."""
    from . import _native
    from .transcripts import TranscriptIDParseError, default_gene_re
    x = Path(x)
    path = x / transcripts
    pat = getattr(parser, "pattern", None)
    is_default = parser is default_parser or (pat is not None and pat.pattern == default_gene_re.pattern and pat.flags == default_gene_re.flags)
    if not is_default or path.stat().st_size == 0:
        return _parse_and_select(out_dir, transcripts, x, top, parser)
    out = Path(out_dir) / (x.stem + "_top.fasta")
    raw = np.memmap(str(path), dtype=np.uint8, mode="r")
    hs, he = _native.fasta_headers(raw)
    if len(hs) == 0:
        return _parse_and_select(out_dir, transcripts, x, top, parser)
    cov, gene, iso, bad = _native.parse_spades_ids(raw, hs, he)
    if bad >= 0:
        fa = Fasta(None, np.zeros(0, np.uint8), np.zeros(len(hs) + 1, np.int64), raw, hs, he)
        raise TranscriptIDParseError(f"Could not parse transcript ID {fa.ids[bad]}.")
    keep = top_gene_mask(cov, gene, top)
    letters, off = _native.fasta_extract(raw, hs, he, keep, alloc)
    starts, lengths = off[:-1], np.diff(off)

    def write():
        _native.fasta_write(out, raw, hs, he, keep, letters, off)

    return out, x.stem, (parser, letters, starts, np.asarray(gene[keep], dtype=np.int32), np.asarray(iso[keep], dtype=np.int32),
                         lengths.astype(np.int32)), write


def _select_and_write(out_dir: Path, transcripts: str, x: Path, *args):
    """Host half of select_top_and_save (thread-safe: no device work): returns (out path, x.stem, records),
    records = (parser, letters, starts, gene, iso, lengths) of what was written."""
    out, stem, rec, write = _parse_and_select(out_dir, transcripts, x, *args)
    write()
    return out, stem, rec


def select_top_and_save(out_dir: Path, transcripts: str, x: Path, *args, preload=None) -> tuple[Path, str]:
    """Writes out_dir/<x.stem>_top.fasta and returns (path, x.stem) (select_top_genes_all.py:12-46).
    preload: a find_homologs.SampleCache that receives the written records right away (the pipeline
    searches them next; the file is then not read back)."""
    out, stem, rec = _select_and_write(out_dir, transcripts, x, *args)
    if preload is not None:
        preload.preload(out, rec[0], rec[1], rec[2], rec[3], rec[4], lengths=rec[5])
    return out, stem


def select_top_and_save_all(out_dir: Path, transcripts: str, dirs, *args, preload=None, jobs: int = 1) -> dict:
    """{top-genes path: sample name} in `dirs` order (filtering_step.py:128-143 runs select_top_and_save over
    the samples with joblib).  The host half (read, parse, select, format, write) runs on up to `jobs`
    threads - the native calls release the GIL - with at most `jobs` + 2 parsed files in memory at a time; the device half (preload) stays on the calling thread, in order."""
    from concurrent.futures import ThreadPoolExecutor
    dirs = [Path(d) for d in dirs]
    workers = max(1, min(int(jobs or 1), len(dirs), 8))
    out = {}
    if workers == 1:
        for d in dirs:
            p, stem = select_top_and_save(out_dir, transcripts, d, *args, preload=preload)
            out[p] = stem
        return out
    with ThreadPoolExecutor(workers) as pool:
        window = workers + 2          # samples submitted but not yet handed to the device
        futs = [pool.submit(_select_and_write, out_dir, transcripts, d, *args) for d in dirs[:window]]
        for k in range(len(dirs)):
            p, stem, rec = futs[k].result()
            futs[k] = None
            if k + window < len(dirs):
                futs.append(pool.submit(_select_and_write, out_dir, transcripts, dirs[k + window], *args))
            if preload is not None:
                preload.preload(p, rec[0], rec[1], rec[2], rec[3], rec[4], lengths=rec[5])
            out[p] = stem
            del rec
    return out
