"""All sample pairs -> gene-matches tables.  Mirrors the reference's find_all_pairs.py:57-233: the
joblib pool of `blastn` jobs over itertools.combinations(inputs, 2) becomes a loop of
rcq_search_pair calls on one CUDA context (every sample is packed + indexed once and stays in HBM:
the role of make_all_dbs / BlastDBCache, :120-159), and each table is labelled and saved exactly as
find_homologs_and_save does (:57-88)."""
from __future__ import annotations

import itertools
import math
import os
from pathlib import Path
from typing import Any, Callable, Iterable, Iterator, Mapping, Optional

import pandas as pd

from .find_homologs import HomologFinder, SampleCache, default_cache
from .gene_matches_tables import default_extension, write_table


def find_homologs_and_save(transcripts1: Path, transcripts2: Path, out_path: Path,
                           hf_args: Optional[Iterable] = None, hf_kwargs: Optional[Mapping[str, Any]] = None
                           ) -> pd.DataFrame:
    finder = HomologFinder(*(hf_args or []), **(hf_kwargs or {}))
    table = finder.get_match_table(transcripts1, transcripts2)
    table["ssample"] = str(transcripts1)
    table["qsample"] = str(transcripts2)
    table[["ssample", "qsample"]] = table[["ssample", "qsample"]].astype("category")
    if out_path is not None:
        write_table(table, out_path)
    return table


def make_output_path(dir_: Path, t1: Path, t2: Path, path_to_sample: Optional[Callable] = None,
                     extension: str = "h5") -> Path:
    ts = [t1, t2]
    if path_to_sample:
        ts = [path_to_sample(t) for t in ts]
    return Path(dir_) / ("{}--{}.{}".format(*ts, extension))


def make_all_dbs(cache_dir, inputs: Iterable[Path], jobs: int = 1, parser=None) -> SampleCache:
    """Pack + index every sample on the device (replaces makeblastdb per sample, :134-159).
    cache_dir is accepted and unused: the 'database' lives in HBM."""
    cache = default_cache()
    for p in inputs:
        cache.get(Path(p), *([parser] if parser is not None else []))
    return cache


def find_all_pairs(inputs: Iterable[Path], output_dir: Path, cache_dir: Path, path_to_sample: Callable[[Path], str],
                   hf_args: Iterable = (), jobs: int = max((os.cpu_count() or 2) - 1, 1)
                   ) -> tuple[Iterator[pd.DataFrame], Iterator[Path], int]:
    """Returns (generator of tables, iterator of table paths in combinations order, C(n, 2)).
    Tables are produced lazily, saved as they are produced.  `jobs` is accepted and unused."""
    inputs = list(inputs)
    hf_args = list(hf_args)
    ext = default_extension()
    cache = make_all_dbs(cache_dir, inputs, parser=hf_args[0] if hf_args else None)

    def tables():
        for t1, t2 in itertools.combinations(inputs, 2):
            yield find_homologs_and_save(t1, t2, make_output_path(output_dir, t1, t2, path_to_sample, ext),
                                         hf_args, {"db_cache": cache})

    paths = (make_output_path(output_dir, t1, t2, path_to_sample, ext) for t1, t2 in itertools.combinations(inputs, 2))
    return tables(), paths, math.comb(len(inputs), 2)
