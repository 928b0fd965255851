"""ctypes binding of the CPU oracle (oracle/rcq_oracle.c).  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the rna_clique_b200 package."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "librcq_oracle.so"

HSP_COLS = ["tx1", "tx2", "minus", "a0", "a1", "b0", "b1", "score2", "mismatch", "gaps", "gapopen", "nident"]
ROW_COLS = ["length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "score", "gaps",
            "nident", "qgene", "qiso", "sgene", "siso", "qlen", "minus", "reverse"]


def build() -> Path:
    src = HERE / "rcq_oracle.c"
    if not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE)], check=True, capture_output=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(LIB))
        L.orc_sample_new.restype = C.c_void_p
        L.orc_sample_new.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        L.orc_sample_free.argtypes = [C.c_void_p]
        L.orc_sample_dust.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set_dust.argtypes = [C.c_int]
        L.orc_search_pair.restype = C.c_void_p
        L.orc_search_pair.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int]
        L.orc_search_pair_directions.restype = C.c_void_p
        L.orc_search_pair_directions.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int]
        L.orc_match_table_from_rows.restype = C.c_void_p
        L.orc_match_table_from_rows.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_int]
        L.orc_result_free.argtypes = [C.c_void_p]
        L.orc_result_counts.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_result_hsps.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_result_rows.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_length_adjustment.restype = C.c_int32
        L.orc_length_adjustment.argtypes = [C.c_int64, C.c_int64, C.c_int64]
        L.orc_evalue.restype = C.c_double
        L.orc_evalue.argtypes = [C.c_int32, C.c_int64, C.c_int64, C.c_int64]
        L.orc_bitscore.restype = C.c_double
        L.orc_bitscore.argtypes = [C.c_int32]
        L.orc_set_gapped_mode.argtypes = [C.c_int]
        L.orc_max_band.restype = C.c_int32
        L.orc_max_band.argtypes = [C.c_int]
        L.orc_extend.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_int32, C.c_int, C.c_void_p]
        _lib = L
    return _lib


class Sample:
    def __init__(self, ascii_: np.ndarray, offsets: np.ndarray, gene: np.ndarray, iso: np.ndarray):
        self.ascii = np.ascontiguousarray(ascii_, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.gene = np.ascontiguousarray(gene, dtype=np.int32)
        self.iso = np.ascontiguousarray(iso, dtype=np.int32)
        self.h = lib().orc_sample_new(self.ascii.ctypes.data, self.offsets.ctypes.data,
                                      self.gene.ctypes.data, self.iso.ctypes.data, len(self.gene))

    def dust(self) -> np.ndarray:
        """Symmetric-DUST mask (level 20, window 64), one byte per base of the concatenated transcripts."""
        out = np.zeros(int(self.offsets[-1] - self.offsets[0]) if len(self.offsets) else 0, dtype=np.uint8)
        lib().orc_sample_dust(self.h, out.ctypes.data)
        return out

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.orc_sample_free(self.h)
            self.h = None


class Result:
    def __init__(self, h):
        self.h = h
        c = np.zeros(7, dtype=np.int64)
        lib().orc_result_counts(h, c.ctypes.data)
        self.counts = dict(zip(["gapped", "fwd", "bwd", "rows", "seeds", "ungapped", "dp_cells"], c.tolist()))

    def hsps(self) -> np.ndarray:
        out = np.zeros((self.counts["gapped"], 12), dtype=np.int32)
        lib().orc_result_hsps(self.h, out.ctypes.data)
        return out

    def rows(self, which: int = 2) -> np.ndarray:
        n = [self.counts["fwd"], self.counts["bwd"], self.counts["rows"]][which]
        out = np.zeros((n, 17), dtype=np.int32)
        lib().orc_result_rows(self.h, which, out.ctypes.data)
        return out

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.orc_result_free(self.h)
            self.h = None


def search_pair(s1: Sample, s2: Sample, evalue: float = 1e-99, top_n: int = 1, keep_all: bool = True,
                independent_directions: bool = False) -> Result:
    """independent_directions: run the two blastn directions of find_homologs.py:229-239 as two searches with
    their own query roles instead of deriving both frames from one search (see rcq_oracle.c)."""
    fn = lib().orc_search_pair_directions if independent_directions else lib().orc_search_pair
    return Result(fn(s1.h, s2.h, evalue, top_n, int(keep_all)))


def match_table_from_rows(fwd: np.ndarray, bwd: np.ndarray, top_n: int = 1, keep_all: bool = True) -> Result:
    fwd = np.ascontiguousarray(fwd, dtype=np.int32).reshape(-1, 17)
    bwd = np.ascontiguousarray(bwd, dtype=np.int32).reshape(-1, 17)
    r = Result.__new__(Result)
    r.h = lib().orc_match_table_from_rows(fwd.ctypes.data, len(fwd), bwd.ctypes.data, len(bwd), top_n, int(keep_all))
    c = np.zeros(7, dtype=np.int64)
    lib().orc_result_counts(r.h, c.ctypes.data)
    r.counts = dict(zip(["gapped", "fwd", "bwd", "rows", "seeds", "ungapped", "dp_cells"], c.tolist()))
    return r


def max_band(reset: bool = True) -> int:
    """Widest wave (in diagonals) any greedy extension has had since the last reset."""
    return int(lib().orc_max_band(int(reset)))


def set_dust(on: bool) -> None:
    """DUST soft-masking of seeds for samples created AFTER the call (default on, as in blastn)."""
    lib().orc_set_dust(int(bool(on)))


def set_gapped_mode(greedy: bool) -> None:
    """True (default): greedy X-drop waves, the canonical gapped stage.  False: the anti-diagonal
    X-drop DP kept as a cross-check (same scores and end points; it splits ties between two
    mismatches and an insertion+deletion pair differently)."""
    lib().orc_set_gapped_mode(int(bool(greedy)))


def extend(q: bytes, s: bytes, greedy: bool = True) -> dict:
    """One gapped extension of ASCII q against ASCII s from (0, 0)."""
    out = np.zeros(6, dtype=np.int32)
    lib().orc_extend(q, len(q), s, len(s), int(bool(greedy)), out.ctypes.data)
    return dict(zip(["score2", "qi", "sj", "mismatch", "gaps", "gapopen"], out.tolist()))
