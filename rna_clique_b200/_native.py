"""ctypes loader for librnaclique_b200.so (the C ABI declared in include/rcq.h).

There is no CPU fallback: if the library is missing, or no CUDA device is present, every
entry point that needs the GPU raises."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("RCQ_LIB") or HERE / "librnaclique_b200.so")   # RCQ_LIB: an experimental build of the same library

TABLE_COLS = ["length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "score", "gaps",
              "nident", "qgene", "qiso", "sgene", "siso", "qlen"]
HSP_COLS = ["tx1", "tx2", "minus", "a0", "a1", "b0", "b1", "score2", "mismatch", "gaps", "gapopen", "nident"]

# every symbol include/rcq.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "rcq_last_error", "rcq_version", "rcq_ctx_create", "rcq_ctx_destroy", "rcq_ctx_sync", "rcq_ctx_stream",
    "rcq_ctx_launches", "rcq_ctx_stage_times", "rcq_top_gene_mask", "rcq_fasta_scan", "rcq_fasta_parse", "rcq_fasta_headers", "rcq_fasta_extract", "rcq_fasta_write", "rcq_parse_spades_ids", "rcq_fasta_format", "rcq_graph_pickle", "rcq_graph_pickle_groups", "rcq_graph_pickle_tables", "rcq_matrix_h5_write", "rcq_sample_create", "rcq_sample_free", "rcq_sample_info", "rcq_sample_download",
    "rcq_ctx_set_dust", "rcq_sample_seed_mask", "rcq_dust_mask_static", "rcq_int32_issue_peak",
    "rcq_search_pair", "rcq_table_rows", "rcq_table_download", "rcq_table_free", "rcq_debug_last_hsps", "rcq_debug_keep_hsps",
    "rcq_length_adjustment", "rcq_table_stats", "rcq_evalue", "rcq_bitscore", "rcq_phase2_host", "rcq_phase2_tables", "rcq_phase2_pack", "rcq_phase2_packed",
    "rcq_table_flat_words", "rcq_table_flatten", "rcq_table_from_flat", "rcq_sample_meta", "rcq_sample_export", "rcq_sample_import",
]


class RcqError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"librnaclique_b200 error {code}: {msg}")
        self.code = code


class NoIdealComponents(RcqError):
    pass


class SearchStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ["n_candidates", "n_seeds", "n_ungapped", "n_gapped", "dp_cells",
                                         "ungapped_cells", "n_fwd", "n_bwd", "n_rows", "band_retries", "n_verified"]]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class TableCols(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in TABLE_COLS] + [("minus", C.c_void_p), ("reverse", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(str(LIB_PATH))
        vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
        L.rcq_last_error.restype = C.c_char_p
        L.rcq_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.rcq_ctx_destroy.argtypes = [vp]
        L.rcq_ctx_destroy.restype = None
        L.rcq_ctx_sync.argtypes = [vp]
        L.rcq_ctx_stream.argtypes = [vp]
        L.rcq_ctx_stream.restype = vp
        L.rcq_ctx_launches.argtypes = [vp]
        L.rcq_ctx_launches.restype = i64
        L.rcq_ctx_stage_times.argtypes = [vp, vp, vp, C.c_int]
        L.rcq_sample_create.argtypes = [vp, vp, i64, C.c_int, vp, vp, vp, vp, i32, C.POINTER(vp)]
        L.rcq_sample_free.argtypes = [vp]
        L.rcq_sample_free.restype = None
        L.rcq_sample_info.argtypes = [vp, vp]
        L.rcq_sample_download.argtypes = [vp, vp, vp, vp, vp, vp]
        L.rcq_ctx_set_dust.argtypes = [vp, C.c_int]
        L.rcq_int32_issue_peak.argtypes = [vp, i32, C.POINTER(dbl)]
        L.rcq_sample_seed_mask.argtypes = [vp, vp, vp]
        L.rcq_dust_mask_static.argtypes = [C.c_int, vp, vp, C.c_int64, vp, vp, vp]
        L.rcq_search_pair.argtypes = [vp, vp, vp, dbl, i32, i32, C.POINTER(vp), C.POINTER(SearchStats)]
        L.rcq_table_rows.argtypes = [vp]
        L.rcq_table_rows.restype = i64
        L.rcq_table_download.argtypes = [vp, C.POINTER(TableCols)]
        L.rcq_table_free.argtypes = [vp]
        L.rcq_table_free.restype = None
        L.rcq_debug_last_hsps.argtypes = [vp, vp, i64]
        L.rcq_debug_last_hsps.restype = i64
        L.rcq_fasta_scan.argtypes = [vp, i64, vp, vp]
        L.rcq_fasta_parse.argtypes = [vp, i64, vp, vp, vp, vp]
        L.rcq_fasta_headers.argtypes = [vp, i64, i64, vp, vp]
        L.rcq_fasta_extract.argtypes = [vp, i64, vp, vp, i64, vp, i64, vp, i64, vp]
        L.rcq_fasta_extract.restype = i64
        L.rcq_fasta_write.argtypes = [C.c_char_p, vp, vp, vp, i64, vp, i64, vp, vp, C.c_int32]
        L.rcq_fasta_write.restype = i64
        L.rcq_parse_spades_ids.argtypes = [vp, vp, vp, i64, vp, vp, vp]
        L.rcq_parse_spades_ids.restype = i64
        L.rcq_fasta_format.argtypes = [vp, vp, vp, i64, vp, vp, C.c_int32, vp]
        L.rcq_graph_pickle.argtypes = [C.c_char_p, C.c_int32, vp, C.c_int32, vp, vp, vp, vp, vp, i64, vp, vp, vp]
        L.rcq_fasta_format.restype = i64
        L.rcq_graph_pickle_groups.argtypes = [C.c_char_p, C.c_int32, vp, C.c_int32, vp, vp, vp, vp, vp, i64, vp, vp, vp]
        L.rcq_graph_pickle_tables.argtypes = [vp, C.c_char_p, C.c_int32, vp, C.c_int32, vp, vp, vp, vp]
        L.rcq_matrix_h5_write.argtypes = [C.c_char_p, C.c_char_p, i64, i64, vp, vp, vp]
        L.rcq_top_gene_mask.argtypes = [vp, vp, i64, i64, vp]
        L.rcq_top_gene_mask.restype = i64
        L.rcq_debug_keep_hsps.argtypes = [vp, C.c_int]
        L.rcq_debug_keep_hsps.restype = C.c_int
        L.rcq_length_adjustment.argtypes = [i64, i64, i64]
        L.rcq_length_adjustment.restype = i32
        L.rcq_table_stats.argtypes = [i64, vp, vp, vp, vp, vp, i64, i64, i64, i64, dbl, dbl, vp, vp, vp]
        L.rcq_evalue.argtypes = [i32, i64, i64, i64]
        L.rcq_evalue.restype = dbl
        L.rcq_bitscore.argtypes = [i32]
        L.rcq_bitscore.restype = dbl
        L.rcq_phase2_host.argtypes = [vp, i32, i64, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp, vp, vp, vp]
        L.rcq_phase2_tables.argtypes = [vp, i32, i64, i32, vp, vp, vp, i32, vp, vp, vp, vp]
        L.rcq_phase2_pack.argtypes = [vp, i32, vp, vp]
        L.rcq_phase2_packed.argtypes = [vp, i32, i64, i32, vp, vp, vp, vp, vp, vp, i32, vp, vp, vp, vp]
        L.rcq_sample_meta.argtypes = [vp, vp]
        L.rcq_sample_export.argtypes = [vp, vp, vp]
        L.rcq_sample_import.argtypes = [vp, vp, vp, C.POINTER(vp)]
        L.rcq_table_flat_words.argtypes = [vp]
        L.rcq_table_flat_words.restype = i64
        L.rcq_table_flatten.argtypes = [vp, vp, vp]
        L.rcq_table_from_flat.argtypes = [vp, vp, i64, C.POINTER(vp)]
        _lib = L
    return _lib


def check(code: int):
    if code != 0:
        msg = lib().rcq_last_error().decode(errors="replace")
        if code == -4:
            raise NoIdealComponents(code, msg)
        raise RcqError(code, msg)


def _ptr(a):
    return None if a is None else a.ctypes.data


class Context:
    """One GPU + one stream (rcq_ctx)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        check(lib().rcq_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device

    def sync(self):
        check(lib().rcq_ctx_sync(self.h))

    @property
    def stream(self) -> int:
        return lib().rcq_ctx_stream(self.h)

    @property
    def launches(self) -> int:
        return lib().rcq_ctx_launches(self.h)

    STAGES = ["pack_index", "seed_join", "ungapped", "sort_group", "gapped_dp", "select_table", "phase2", "seed_probe"]

    def stage_times(self, reset: bool = False) -> dict:
        ms = np.zeros(8, dtype=np.float64)
        calls = np.zeros(8, dtype=np.int64)
        check(lib().rcq_ctx_stage_times(self.h, ms.ctypes.data, calls.ctypes.data, int(reset)))
        return {n: dict(ms=float(ms[i]), calls=int(calls[i])) for i, n in enumerate(self.STAGES)}

    def set_dust(self, on: bool) -> bool:
        """DUST soft-masking of seeds for samples created on this context from now on (default on, as in
        blastn); returns the previous setting."""
        return bool(lib().rcq_ctx_set_dust(self.h, int(bool(on))))

    def int32_issue_peak(self, reps: int = 5) -> float:
        """Measured INT32 issue ceiling of the device, lane-operations per second (rcq_int32_issue_peak)."""
        v = C.c_double()
        check(lib().rcq_int32_issue_peak(self.h, int(reps), C.byref(v)))
        return float(v.value)

    def close(self):
        if self.h:
            lib().rcq_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Sample:
    """Device-resident packed sample + seed index (rcq_sample)."""

    def __init__(self, ctx: Context, ascii_, offsets, gene, iso, name: str = "", lengths=None):
        """ascii_: uint8 numpy array or torch tensor (CPU or CUDA).  offsets: int64 [n+1] boundaries of
        contiguous transcripts, or — when `lengths` is given — int64 [n] start positions."""
        self.ctx = ctx
        self.name = name
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        gene = np.ascontiguousarray(gene, dtype=np.int32)
        iso = np.ascontiguousarray(iso, dtype=np.int32)
        n = len(gene)
        if lengths is None:
            assert len(offsets) == n + 1
            starts = np.ascontiguousarray(offsets[:-1])
            lengths = np.diff(offsets).astype(np.int32)
        else:
            starts = offsets
            lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        assert len(starts) == n and len(lengths) == n and len(iso) == n
        on_device = 0
        if hasattr(ascii_, "data_ptr"):  # torch tensor
            alen = int(ascii_.numel())
            if ascii_.is_cuda:
                on_device, p = 1, ascii_.data_ptr()
                keep = ascii_
            else:
                keep = ascii_.numpy()
                p = keep.ctypes.data
        else:
            keep = np.ascontiguousarray(ascii_, dtype=np.uint8)
            p, alen = keep.ctypes.data, len(keep)
        h = C.c_void_p()
        check(lib().rcq_sample_create(ctx.h, p, alen, on_device, starts.ctypes.data, lengths.ctypes.data,
                                      gene.ctypes.data, iso.ctypes.data, n, C.byref(h)))
        del keep
        self.h = h
        self._keep = None
        info = np.zeros(5, dtype=np.int64)
        check(lib().rcq_sample_info(self.h, info.ctypes.data))
        self.n_seq, self.total, self.arena, self.n_index, self.n_masked = (int(x) for x in info)

    def meta(self) -> np.ndarray:
        m = np.zeros(8, dtype=np.int64)
        check(lib().rcq_sample_meta(self.h, m.ctypes.data))
        return m

    def export_flat(self, dev_ptr: int) -> None:
        """Copies every device array of the sample into the flat device buffer at dev_ptr (meta()[7] bytes)."""
        check(lib().rcq_sample_export(self.ctx.h, self.h, dev_ptr))

    @classmethod
    def import_flat(cls, ctx: "Context", meta: np.ndarray, dev_ptr: int, name: str = "") -> "Sample":
        meta = np.ascontiguousarray(meta, dtype=np.int64)
        h = C.c_void_p()
        check(lib().rcq_sample_import(ctx.h, meta.ctypes.data, dev_ptr, C.byref(h)))
        s = cls.__new__(cls)
        s.ctx, s.name, s.h, s._keep = ctx, name, h, None
        s.n_seq, s.total, s.arena, s.n_index, s.n_masked = (int(x) for x in meta[:5])
        return s

    def seed_mask(self):
        """(seeding mask words uint32[arena/32], bases masked by DUST alone)."""
        m = np.zeros(self.arena // 32, dtype=np.uint32)
        nd = np.zeros(1, dtype=np.int64)
        check(lib().rcq_sample_seed_mask(self.h, m.ctypes.data, nd.ctypes.data))
        return m, int(nd[0])

    def download(self):
        fwd = np.zeros(self.arena // 16, dtype=np.uint32)
        rc = np.zeros(self.arena // 16, dtype=np.uint32)
        mask = np.zeros(self.arena // 32, dtype=np.uint32)
        off = np.zeros(self.n_seq, dtype=np.int64)
        index = np.zeros(self.n_index, dtype=np.uint64)
        check(lib().rcq_sample_download(self.h, fwd.ctypes.data, rc.ctypes.data, mask.ctypes.data,
                                        off.ctypes.data, index.ctypes.data))
        return dict(fwd=fwd, rc=rc, mask=mask, off=off, index=index)

    def close(self):
        if getattr(self, "h", None) and self.ctx.h:
            lib().rcq_sample_free(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Table:
    """Device-resident gene-matches table (rcq_table)."""

    def __init__(self, ctx: Context, h, stats=None):
        self.ctx, self.h, self.stats = ctx, h, stats

    def __len__(self):
        return int(lib().rcq_table_rows(self.h))

    def download(self) -> dict:
        n = len(self)
        block = np.empty((len(TABLE_COLS), n), dtype=np.int32)       # one allocation for the fifteen int32 columns
        cols = {c: block[k] for k, c in enumerate(TABLE_COLS)}
        cols["minus"] = np.empty(n, dtype=np.uint8)
        cols["reverse"] = np.empty(n, dtype=np.uint8)
        tc = TableCols(**{k: v.ctypes.data for k, v in cols.items()})
        check(lib().rcq_table_download(self.h, C.byref(tc)))
        return cols

    def close(self):
        if getattr(self, "h", None) and self.ctx.h:
            lib().rcq_table_free(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def search_pair(ctx: Context, s1: Sample, s2: Sample, evalue: float = 1e-99, top_n: int = 1,
                keep_all: bool = True) -> Table:
    h = C.c_void_p()
    st = SearchStats()
    check(lib().rcq_search_pair(ctx.h, s1.h, s2.h, float(evalue), int(top_n), int(bool(keep_all)),
                                C.byref(h), C.byref(st)))
    return Table(ctx, h, st.as_dict())


def fasta_parse(raw: np.ndarray):
    """(ascii uint8[n_letters], offsets int64[n+1], hdr_start int64[n], hdr_end int64[n]) of FASTA bytes."""
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    cnt = np.zeros(2, dtype=np.int64)
    check(lib().rcq_fasta_scan(raw.ctypes.data, len(raw), cnt.ctypes.data, cnt.ctypes.data + 8))
    n, letters = int(cnt[0]), int(cnt[1])      # letters: an upper bound (line ends are still counted)
    ascii_ = np.empty(letters, dtype=np.uint8)
    offsets = np.zeros(n + 1, dtype=np.int64)
    hs, he = np.zeros(n, dtype=np.int64), np.zeros(n, dtype=np.int64)
    check(lib().rcq_fasta_parse(raw.ctypes.data, len(raw), ascii_.ctypes.data, offsets.ctypes.data, hs.ctypes.data, he.ctypes.data))
    ascii_ = ascii_[:int(offsets[n])]
    return ascii_, offsets, hs, he


def fasta_headers(raw: np.ndarray):
    """(hdr_start int64[n], hdr_end int64[n]) of FASTA bytes: the header line of every record, sequence bytes untouched."""
    cnt = np.zeros(2, dtype=np.int64)
    check(lib().rcq_fasta_scan(raw.ctypes.data, len(raw), cnt.ctypes.data, cnt.ctypes.data + 8))
    n = int(cnt[0])
    hs, he = np.zeros(n, dtype=np.int64), np.zeros(n, dtype=np.int64)
    check(lib().rcq_fasta_headers(raw.ctypes.data, len(raw), n, hs.ctypes.data, he.ctypes.data))
    return hs, he


def fasta_extract(raw: np.ndarray, hs: np.ndarray, he: np.ndarray, keep: np.ndarray, alloc=None):
    """Letters of the records `keep` (ascending record numbers): (uint8 buffer holding them, offsets int64[len(keep)+1]).
    alloc(nbytes) -> uint8 array of at least that size (e.g. pinned memory); numpy.empty by default."""
    keep = np.ascontiguousarray(keep, dtype=np.int64)
    nxt = np.append(hs[1:], len(raw))
    bound = int((nxt[keep] - he[keep]).sum()) if len(keep) else 0
    out = (alloc or (lambda b: np.empty(b, dtype=np.uint8)))(max(bound, 1))
    off = np.zeros(len(keep) + 1, dtype=np.int64)
    L = lib()
    tot = L.rcq_fasta_extract(raw.ctypes.data, len(raw), hs.ctypes.data, he.ctypes.data, len(hs), keep.ctypes.data, len(keep),
                              out.ctypes.data, len(out), off.ctypes.data)
    if tot < 0:
        raise RuntimeError(L.rcq_last_error().decode())
    return out[:max(int(tot), 1)], off


def fasta_write(path, raw: np.ndarray, hs: np.ndarray, he: np.ndarray, keep: np.ndarray, letters: np.ndarray, off: np.ndarray,
                width: int = 60) -> int:
    """rcq_fasta_write: the records `keep` as FASTA text (headers from the file bytes, letters from fasta_extract)."""
    keep = np.ascontiguousarray(keep, dtype=np.int64)
    off = np.ascontiguousarray(off, dtype=np.int64)
    L = lib()
    n = L.rcq_fasta_write(str(path).encode(), raw.ctypes.data, hs.ctypes.data, he.ctypes.data, len(hs), keep.ctypes.data, len(keep),
                          letters.ctypes.data, off.ctypes.data, int(width))
    if n < 0:
        raise RuntimeError(L.rcq_last_error().decode())
    return int(n)


def fasta_format(ascii_: np.ndarray, starts, lengths, headers, width: int = 60) -> np.ndarray:
    """The FASTA text (uint8 array) of the given records."""
    ascii_ = np.ascontiguousarray(ascii_, dtype=np.uint8)
    starts = np.ascontiguousarray(starts, dtype=np.int64)
    lengths = np.ascontiguousarray(lengths, dtype=np.int64)
    if isinstance(headers, tuple):          # (uint8 text, int64 offsets), Fasta.header_bytes
        hdr = np.ascontiguousarray(headers[0], dtype=np.uint8)
        hdr_off = np.ascontiguousarray(headers[1], dtype=np.int64)
        hb = range(len(hdr_off) - 1)
    else:
        hb = [h.encode() for h in headers]
        hdr = np.frombuffer(b"".join(hb), dtype=np.uint8) if hb else np.zeros(0, np.uint8)
        hdr_off = np.zeros(len(hb) + 1, dtype=np.int64)
        np.cumsum([len(b) for b in hb], out=hdr_off[1:])
    args = (ascii_.ctypes.data, starts.ctypes.data, lengths.ctypes.data, len(hb), hdr.ctypes.data if len(hdr) else None,
            hdr_off.ctypes.data, int(width))
    size = lib().rcq_fasta_format(*args, None)
    if size < 0:
        check(int(size))
    out = np.empty(size, dtype=np.uint8)
    lib().rcq_fasta_format(*args, out.ctypes.data)
    return out


def graph_pickle(path, sample_names, groups, isolated=()) -> dict:
    """rcq_graph_pickle: groups = [(s index, q index, sgene int32[], qgene int32[])], isolated = [(s index, gene)].
    Writes the networkx.Graph pickle of the gene-matches graph; returns dict(nodes, edges, bytes)."""
    names = [str(n).encode() for n in sample_names]
    arr = (C.c_char_p * max(len(names), 1))(*names)
    g_s = np.ascontiguousarray([g[0] for g in groups], dtype=np.int32)
    g_q = np.ascontiguousarray([g[1] for g in groups], dtype=np.int32)
    cols = [(np.ascontiguousarray(g[2], dtype=np.int32), np.ascontiguousarray(g[3], dtype=np.int32)) for g in groups]
    g_len = np.ascontiguousarray([len(c[0]) for c in cols], dtype=np.int64)
    sp = (C.c_void_p * max(len(cols), 1))(*[c[0].ctypes.data for c in cols])
    qp = (C.c_void_p * max(len(cols), 1))(*[c[1].ctypes.data for c in cols])
    iso_s = np.ascontiguousarray([i[0] for i in isolated], dtype=np.int32)
    iso_g = np.ascontiguousarray([i[1] for i in isolated], dtype=np.int32)
    counts = np.zeros(3, dtype=np.int64)
    check(lib().rcq_graph_pickle_groups(str(path).encode(), len(names), C.cast(arr, C.c_void_p), len(groups), _ptr(g_s), _ptr(g_q),
                                        _ptr(g_len), C.cast(sp, C.c_void_p), C.cast(qp, C.c_void_p), len(iso_s), _ptr(iso_s),
                                        _ptr(iso_g), counts.ctypes.data))
    return dict(nodes=int(counts[0]), edges=int(counts[1]), bytes=int(counts[2]))


def graph_pickle_tables(ctx: "Context", path, sample_names, tables, t_s, t_q) -> dict:
    """rcq_graph_pickle_tables: graph.pkl from device-resident tables (tables[k] joins samples t_s[k], t_q[k]).
    Raises NativeError (RCQ_ERR_ARG / RCQ_ERR_CAPACITY) when the tables do not qualify for the device path."""
    names = [str(n).encode() for n in sample_names]
    arr = (C.c_char_p * max(len(names), 1))(*names)
    hp = (C.c_void_p * max(len(tables), 1))(*[t.h for t in tables])
    ts = np.ascontiguousarray(t_s, dtype=np.int32)
    tq = np.ascontiguousarray(t_q, dtype=np.int32)
    counts = np.zeros(3, dtype=np.int64)
    check(lib().rcq_graph_pickle_tables(ctx.h, str(path).encode(), len(names), C.cast(arr, C.c_void_p), len(tables),
                                        C.cast(hp, C.c_void_p), _ptr(ts), _ptr(tq), counts.ctypes.data))
    return dict(nodes=int(counts[0]), edges=int(counts[1]), bytes=int(counts[2]))


def parse_spades_ids(raw: np.ndarray, hdr_start: np.ndarray, hdr_end: np.ndarray):
    """(coverage, gene, iso, first_bad): the default rnaSPAdes id pattern over every header; first_bad = -1
    when all ids matched."""
    n = len(hdr_start)
    cov, gene, iso = np.zeros(n, dtype=np.float64), np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
    bad = lib().rcq_parse_spades_ids(raw.ctypes.data, hdr_start.ctypes.data, hdr_end.ctypes.data, n,
                                     cov.ctypes.data, gene.ctypes.data, iso.ctypes.data)
    return cov, gene, iso, int(bad)


def top_gene_mask(coverage, gene, top: int) -> np.ndarray:
    """Indices (file order) of the transcripts whose gene is among the `top` genes (rcq_top_gene_mask)."""
    coverage = np.ascontiguousarray(coverage, dtype=np.float64)
    gene = np.ascontiguousarray(gene, dtype=np.int32)
    keep = np.zeros(len(gene), dtype=np.uint8)
    r = lib().rcq_top_gene_mask(coverage.ctypes.data, gene.ctypes.data, len(gene), int(top), keep.ctypes.data)
    if r < 0:
        check(int(r))
    return np.flatnonzero(keep)


def debug_keep_hsps(ctx: Context, on: bool = True) -> bool:
    """Switches the per-search copy of the gapped HSP set on or off; returns the previous setting."""
    return bool(lib().rcq_debug_keep_hsps(ctx.h, int(on)))


def debug_last_hsps(ctx: Context) -> np.ndarray:
    n = lib().rcq_debug_last_hsps(ctx.h, None, 0)
    out = np.zeros((n, 12), dtype=np.int32)
    if n:
        lib().rcq_debug_last_hsps(ctx.h, out.ctypes.data, n)
    return out


def phase2_host(ctx: Context, n_samples: int, stride: int, row_off, t_s, t_q, sgene, qgene, nident, length,
                gaps, sample_count=None, want_valid: bool = True, t_flags=None):
    """rcq_phase2_host: concatenated host tables -> (nident_sum, denom_sum, valid, counts).
    Raises NoIdealComponents (after filling the sums) like the reference's NoIdealComponentsError."""
    row_off = np.ascontiguousarray(row_off, dtype=np.int64)
    n_tables = len(row_off) - 1
    arrs = [np.ascontiguousarray(a, dtype=np.int32) for a in (t_s, t_q, sgene, qgene, nident, length, gaps)]
    ni = np.zeros(max(n_tables, 1), dtype=np.int64)
    de = np.zeros(max(n_tables, 1), dtype=np.int64)
    valid = np.zeros(n_samples * stride, dtype=np.uint8) if want_valid else None
    counts = np.zeros(6, dtype=np.int64)
    if t_flags is not None:
        t_flags = np.ascontiguousarray(t_flags, dtype=np.uint8)
        assert len(t_flags) == n_tables
    code = lib().rcq_phase2_host(ctx.h, n_samples, stride, n_tables, row_off.ctypes.data, *[a.ctypes.data for a in arrs],
                                 _ptr(t_flags), int(sample_count or 0), ni.ctypes.data, de.ctypes.data, _ptr(valid), counts.ctypes.data)
    res = dict(nident_sum=ni[:n_tables], denom_sum=de[:n_tables], valid=valid,
               counts=dict(zip(["components", "ideal", "sample_count", "nodes", "edges", "ge_samples"], counts.tolist())))
    if code == -4:
        res["no_ideal"] = True
        return res
    check(code)
    res["no_ideal"] = False
    return res


def phase2_tables(ctx: Context, n_samples: int, stride: int, tables, t_s, t_q, sample_count=None,
                  want_valid: bool = True):
    """rcq_phase2_tables: device-resident tables straight from search_pair."""
    n_tables = len(tables)
    hs = (C.c_void_p * max(n_tables, 1))(*[t.h for t in tables])
    t_s = np.ascontiguousarray(t_s, dtype=np.int32)
    t_q = np.ascontiguousarray(t_q, dtype=np.int32)
    ni = np.zeros(max(n_tables, 1), dtype=np.int64)
    de = np.zeros(max(n_tables, 1), dtype=np.int64)
    valid = np.zeros(n_samples * stride, dtype=np.uint8) if want_valid else None
    counts = np.zeros(6, dtype=np.int64)
    code = lib().rcq_phase2_tables(ctx.h, n_samples, stride, n_tables, hs, t_s.ctypes.data, t_q.ctypes.data,
                                   int(sample_count or 0), ni.ctypes.data, de.ctypes.data, _ptr(valid), counts.ctypes.data)
    res = dict(nident_sum=ni[:n_tables], denom_sum=de[:n_tables], valid=valid,
               counts=dict(zip(["components", "ideal", "sample_count", "nodes", "edges", "ge_samples"], counts.tolist())))
    if code == -4:
        res["no_ideal"] = True
        return res
    check(code)
    res["no_ideal"] = False
    return res


def phase2_pack(ctx: Context, tables, dev_ptr: int) -> None:
    """rcq_phase2_pack: the five phase-2 columns of `tables` as one column-major int32 block at dev_ptr."""
    hs = (C.c_void_p * max(len(tables), 1))(*[t.h for t in tables])
    check(lib().rcq_phase2_pack(ctx.h, len(tables), hs, dev_ptr))


def phase2_packed(ctx: Context, n_samples: int, stride: int, row_off, t_s, t_q, src_base, src_stride, dev_ptr: int,
                  sample_count=None, want_valid: bool = False):
    """rcq_phase2_packed: phase 2 on the all-gathered five-column blocks."""
    row_off = np.ascontiguousarray(row_off, dtype=np.int64)
    n_tables = len(row_off) - 1
    t_s = np.ascontiguousarray(t_s, dtype=np.int32); t_q = np.ascontiguousarray(t_q, dtype=np.int32)
    src_base = np.ascontiguousarray(src_base, dtype=np.int64); src_stride = np.ascontiguousarray(src_stride, dtype=np.int64)
    ni = np.zeros(max(n_tables, 1), dtype=np.int64)
    de = np.zeros(max(n_tables, 1), dtype=np.int64)
    valid = np.zeros(n_samples * stride, dtype=np.uint8) if want_valid else None
    counts = np.zeros(6, dtype=np.int64)
    code = lib().rcq_phase2_packed(ctx.h, n_samples, stride, n_tables, row_off.ctypes.data, t_s.ctypes.data, t_q.ctypes.data,
                                   src_base.ctypes.data, src_stride.ctypes.data, dev_ptr, int(sample_count or 0),
                                   ni.ctypes.data, de.ctypes.data, _ptr(valid), counts.ctypes.data)
    res = dict(nident_sum=ni[:n_tables], denom_sum=de[:n_tables], valid=valid,
               counts=dict(zip(["components", "ideal", "sample_count", "nodes", "edges", "ge_samples"], counts.tolist())))
    if code == -4:
        res["no_ideal"] = True
        return res
    check(code)
    res["no_ideal"] = False
    return res
