"""End-to-end entry point.  Mirrors the reference's rna_clique.py:28-178: same signature, same
outputs (od1/*_top.fasta, od2/<S1>--<S2>.<ext>, graph pickle, distance matrix), same return value
(SampleSimilarity, {top-genes path: sample name}).  Directories are not created here, as upstream.

matrix file: `.h5` -> the pandas fixed-format frame under the key "matrix" that upstream's
DataFrame.to_hdf(key="matrix") writes, produced by the library's own HDF5 writer (rcq_matrix_h5_write: no
PyTables needed; pd.read_hdf reads it where PyTables exists, h5fixed.read_frame anywhere); any other suffix is
chosen from {.pkl, .csv, .npy}."""
from __future__ import annotations

import os
from pathlib import Path
from typing import Callable, Iterable, Optional

import numpy as np

from . import filtering_step as _fs
from .filtered_distance import SampleSimilarity
from .transcripts import default_parser


def write_matrix(mat, path: Path) -> Path:
    path = Path(path)
    if path.suffix == ".h5":
        from .h5fixed import write_frame
        write_frame(path, "matrix", mat)
    elif path.suffix == ".csv":
        mat.to_csv(path)
    elif path.suffix == ".npy":
        np.save(path, mat.to_numpy())
    else:
        mat.to_pickle(path)
    return path


def rna_clique(dirs: Iterable[Path], out_dir_1: Path, out_dir_2: Path, cache_dir: Path, output_graph: Path,
               output_matrix: Optional[Path], top_genes: int, transcripts: str = "transcripts.fasta",
               top_matches: int = 1, id_parser: Callable = default_parser, evalue: float = 1e-99,
               keep_all: bool = True, store_dfs: bool = False,
               jobs: int = max((os.cpu_count() or 2) - 1, 1)):
    import time
    t0 = time.perf_counter()
    tables, table_paths, graph, num_tables, pts = _fs.filtering_step(
        dirs, out_dir_1, out_dir_2, cache_dir, output_graph, top_genes, transcripts, top_matches, id_parser,
        evalue, keep_all, jobs)
    tables = SampleSimilarity.mapping_from_dfs(tables)
    if store_dfs:
        tables = dict(tables)
    sim = SampleSimilarity(graph, tables, _device=getattr(graph, "__dict__", {}).get("_rcq_device"))
    t1 = time.perf_counter()
    if output_matrix is not None:
        write_matrix(sim.get_dissimilarity_df(), output_matrix)
    if os.environ.get("RCQ_TRACE_API"):
        import sys
        print(f"[rna_clique] filtering_step {t1 - t0:.3f} s, phase 2 + matrix {time.perf_counter() - t1:.3f} s", file=sys.stderr)
    return sim, pts
