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

import numpy as np
import pandas as pd

from .find_homologs import HomologFinder, SampleCache, default_cache
from .gene_matches_tables import default_extension, write_table


def find_homologs_and_save(transcripts1: Path, transcripts2: Path, out_path: Path,
                           hf_args: Optional[Iterable] = None, hf_kwargs: Optional[Mapping[str, Any]] = None
                           ) -> pd.DataFrame:
    finder = HomologFinder(*(hf_args or []), **(hf_kwargs or {}))
    return _label_and_save(finder.get_match_table(transcripts1, transcripts2), transcripts1, transcripts2, out_path)


def _label_and_save(table: pd.DataFrame, transcripts1: Path, transcripts2: Path, out_path: Optional[Path]) -> pd.DataFrame:
    # the same two category columns as assigning the strings and .astype("category") (find_all_pairs.py:82-86),
    # built from codes instead of 2 x len(table) Python strings
    codes = np.zeros(len(table), dtype=np.int8)
    table["ssample"] = pd.Categorical.from_codes(codes, categories=[str(transcripts1)])
    table["qsample"] = pd.Categorical.from_codes(codes, categories=[str(transcripts2)])
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
        # one pair ahead: a helper thread runs the search and the device->host copy of pair k+1 (native
        # calls, GIL released; it is the only user of the cache's context meanwhile) while this thread turns
        # pair k into its DataFrame and saves it
        from concurrent.futures import ThreadPoolExecutor
        from .find_homologs import table_to_frame
        pairs = list(itertools.combinations(inputs, 2))
        finder = HomologFinder(*hf_args, db_cache=cache)

        def device_part(t1, t2):
            t = finder.get_device_table(t1, t2)
            try:
                return t.download(), t.db
            finally:
                t.close()

        with ThreadPoolExecutor(1) as pool:
            fut = pool.submit(device_part, *pairs[0]) if pairs else None
            for k, (t1, t2) in enumerate(pairs):
                cols, db = fut.result()
                fut = pool.submit(device_part, *pairs[k + 1]) if k + 1 < len(pairs) else None
                yield _label_and_save(table_to_frame(cols, *db), t1, t2, make_output_path(output_dir, t1, t2, path_to_sample, ext))

    paths = (make_output_path(output_dir, t1, t2, path_to_sample, ext) for t1, t2 in itertools.combinations(inputs, 2))
    return tables(), paths, math.comb(len(inputs), 2)
