"""Phase 1 driver.  Mirrors the reference's filtering_step.py:61-164: top genes per sample ->
all-pairs tables -> gene-matches graph -> graph pickle; returns lazy table readers."""
from __future__ import annotations

import itertools
import os
import pickle
from pathlib import Path
from typing import Callable, Iterable

from .build_graph import build_graph, dump_graph
from .filtered_distance import SampleSimilarity
from .find_all_pairs import find_all_pairs
from .select_top_genes import select_top_and_save, select_top_and_save_all
from .transcripts import default_parser


def filtering_step(dirs: Iterable[Path], out_dir_1: Path, out_dir_2: Path, cache_dir: Path, output_graph: Path,
                   top_genes: int, transcripts: str = "transcripts.fasta", top_matches: int = 1,
                   id_parser: Callable = default_parser, evalue: float = 1e-99, keep_all: bool = True,
                   jobs: int = max((os.cpu_count() or 2) - 1, 1)):
    # the selected records go to the device as they are written (the role of make_all_dbs right after):
    # od1/*_top.fasta is not read back
    from .find_homologs import default_cache
    path_to_sample = select_top_and_save_all(Path(out_dir_1), transcripts, dirs, top_genes, id_parser,
                                             preload=default_cache(), jobs=jobs)
    tables, table_paths, num_tables = find_all_pairs(path_to_sample, out_dir_2, cache_dir, path_to_sample.__getitem__,
                                                     hf_args=[id_parser, top_matches, evalue, keep_all], jobs=jobs)
    graph = build_graph(tables)
    if output_graph is not None:
        dump_graph(graph, output_graph)
    p1, p2 = itertools.tee(table_paths)
    return map(SampleSimilarity._read_table, p1), p2, graph, num_tables, path_to_sample
