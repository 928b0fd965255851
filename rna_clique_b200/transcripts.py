"""FASTA-id parsing.  Mirrors the reference's transcripts.py:8-130: a transcript id is decoded into
TranscriptID(coverage: float, gene: int, isoform: int) by a regular expression whose groups are
named after the fields or positional in that order; a non-matching id raises
TranscriptIDParseError (transcripts.py:66,110-124)."""
from __future__ import annotations

import re
from typing import Callable, NamedTuple

import numpy as np

# rnaSPAdes ids, e.g. NODE_1_length_15273_cov_210.5_g0_i0 (reference transcripts.py:8; the coverage
# needs a decimal point)
default_gene_re = re.compile(r"^.*cov_([0-9]+(?:\.[0-9]+))_g([0-9]+)_i([0-9]+)")


class TranscriptIDParseError(Exception):
    pass


class _TranscriptIDBase(NamedTuple):
    coverage: float
    gene: int
    isoform: int


class TranscriptID(_TranscriptIDBase):
    """Converts its attributes on construction (reference cast_namedtuple, transcripts.py:10-47)."""

    def __new__(cls, coverage, gene, isoform):
        return super().__new__(cls, float(coverage), int(gene), int(isoform))

    @classmethod
    def parser_from_re(cls, expr: re.Pattern) -> Callable[[str], "TranscriptID"]:
        fields = cls._fields
        named = set(expr.groupindex.values())
        positional = [i for i in range(1, expr.groups + 1) if i not in named]

        def parse_transcript_id(id_: str) -> TranscriptID:
            m = expr.search(id_)
            if m is None:
                raise TranscriptIDParseError(f"Could not parse transcript ID {id_}.")
            d, free = {}, iter(positional)
            for f in fields:
                if f in expr.groupindex:
                    d[f] = m.group(f)
                else:
                    try:
                        d[f] = m.group(next(free))
                    except StopIteration:
                        raise TranscriptIDParseError(f"Could not parse transcript ID {id_}.") from None
            return cls(**d)

        parse_transcript_id.pattern = expr
        return parse_transcript_id


default_parser = TranscriptID.parser_from_re(default_gene_re)


def parse_ids(ids, parser=default_parser):
    """(coverage float64[n], gene int32[n], isoform int32[n]) of a list of FASTA ids."""
    n = len(ids)
    cov = np.empty(n, dtype=np.float64)
    gene = np.empty(n, dtype=np.int32)
    iso = np.empty(n, dtype=np.int32)
    for k, s in enumerate(ids):
        t = parser(s)
        cov[k], gene[k], iso[k] = t.coverage, t.gene, t.isoform
    return cov, gene, iso


def parse_fasta_ids(fa, parser=default_parser):
    """parse_ids over the records of a Fasta.  The default rnaSPAdes pattern is decoded natively from the
    file bytes (rcq_parse_spades_ids, same re.search semantics: the last position that matches wins);
    any other parser runs in Python."""
    pat = getattr(parser, "pattern", None)
    is_default = parser is default_parser or (pat is not None and pat.pattern == default_gene_re.pattern and pat.flags == default_gene_re.flags)
    if is_default and getattr(fa, "raw", None) is not None and len(fa):
        from . import _native
        cov, gene, iso, bad = _native.parse_spades_ids(fa.raw, fa.hdr_start, fa.hdr_end)
        if bad >= 0:
            raise TranscriptIDParseError(f"Could not parse transcript ID {fa.ids[bad]}.")
        return cov, gene, iso
    return parse_ids(fa.ids, parser)
