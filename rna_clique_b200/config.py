"""The analysis configuration YAML of the reference, as far as the hot path reads it.

Mirrors `RNACliqueConfig.yaml_load / yaml_save` (reference config.py:128-169): the same keys, unknown
keys rejected with `YAMLConfigFileError`, `None` values left out on save, paths and the id regex stored
as strings.  Settings missing from the file are derived from `output_dir` with the reference's default
names (config.py:1472-1483: od1, od2, db_cache, graph.pkl, distance_matrix.<ext>, config.yaml).
`run_analysis(cfg)` is what the `rna-clique` command does after parsing (rna_clique.py:180-217): make the
output directories, run `rna_clique()`, save the matrix and the finished config.  The argument-manager
machinery, config chaining and every other sub-command are out of scope.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field, fields
from datetime import datetime
from pathlib import Path
from typing import Optional

import yaml

from .transcripts import TranscriptID, default_gene_re

CONFIG_VERSION = "0.0.1"
DEFAULT_OUT_DIRS = {"top_genes_dir": "od1", "tables_dir": "od2", "cache_dir": "db_cache"}
_PATH_FIELDS = ("top_genes_dir", "tables_dir", "cache_dir", "output_dir", "graph", "matrix", "subset_of")


class YAMLConfigFileError(Exception):
    pass


@dataclass
class RNACliqueConfig:
    config_version: str = CONFIG_VERSION
    title: Optional[str] = None
    input_dirs: Optional[list] = None
    top_genes_dir: Optional[Path] = None
    tables_dir: Optional[Path] = None
    cache_dir: Optional[Path] = None
    output_dir: Optional[Path] = None
    graph: Optional[Path] = None
    top_genes: Optional[int] = None
    transcripts_name: Optional[str] = "transcripts.fasta"
    top_matches: Optional[int] = 1
    evalue: Optional[float] = 1e-99
    keep_all: Optional[bool] = True
    jobs: Optional[int] = None
    transcript_id_regex: Optional[re.Pattern] = field(default_factory=lambda: default_gene_re)
    path_to_sample: Optional[dict] = None
    matrix: Optional[Path] = None
    finished: Optional[datetime] = None
    version: str = "b200"
    subset_of: Optional[Path] = None
    verbose: Optional[bool] = False

    # ---- YAML
    @classmethod
    def from_mapping(cls, m: dict) -> "RNACliqueConfig":
        kw = dict(m)
        for k in _PATH_FIELDS:
            if kw.get(k) is not None:
                kw[k] = Path(kw[k])
        if kw.get("input_dirs") is not None:
            kw["input_dirs"] = [Path(p) for p in kw["input_dirs"]]
        if kw.get("transcript_id_regex") is not None and not isinstance(kw["transcript_id_regex"], re.Pattern):
            kw["transcript_id_regex"] = re.compile(kw["transcript_id_regex"])
        if kw.get("path_to_sample") is not None:
            kw["path_to_sample"] = {Path(p): s for p, s in kw["path_to_sample"].items()}
        if isinstance(kw.get("finished"), str):
            kw["finished"] = datetime.fromisoformat(kw["finished"])
        return cls(**kw)

    @classmethod
    def yaml_load(cls, path: Path) -> "RNACliqueConfig":
        with open(path, "r") as f:
            content = yaml.safe_load(f)
        if not isinstance(content, dict):
            raise YAMLConfigFileError(f"YAML file {path} does not contain a mapping.")
        known = {f.name for f in fields(cls)}
        for key in content:
            if key not in known:
                raise YAMLConfigFileError(f"Unknown configuration setting {key} in {path}.")
        try:
            return cls.from_mapping(content)
        except (TypeError, ValueError, re.error) as e:
            raise YAMLConfigFileError(f"Failed to load YAML config {path}.") from e

    def marshal(self, hide_none: bool = True) -> dict:
        out = {}
        for f in fields(self):
            v = getattr(self, f.name)
            if v is None and hide_none:
                continue
            if isinstance(v, Path):
                v = str(v)
            elif isinstance(v, re.Pattern):
                v = v.pattern
            elif isinstance(v, datetime):
                v = v.isoformat()
            elif f.name == "input_dirs" and v is not None:
                v = [str(p) for p in v]
            elif f.name == "path_to_sample" and v is not None:
                v = {str(p): s for p, s in v.items()}
            out[f.name] = v
        return out

    def yaml_save(self, path: Path) -> None:
        with open(path, "w") as f:
            f.write(yaml.safe_dump(self.marshal(hide_none=True), sort_keys=False))

    def mark_finish(self) -> None:
        self.finished = datetime.now()

    # ---- defaults derived from output_dir (config.py:1472-1483)
    def with_defaults(self) -> "RNACliqueConfig":
        from .gene_matches_tables import default_extension
        if self.output_dir is not None:
            for name, sub in DEFAULT_OUT_DIRS.items():
                if getattr(self, name) is None:
                    setattr(self, name, self.output_dir / sub)
            if self.graph is None:
                self.graph = self.output_dir / "graph.pkl"
            if self.matrix is None:
                # the upstream name (config.py:1482); written by the library's own HDF5 writer (h5fixed.py)
                self.matrix = self.output_dir / "distance_matrix.h5"
        return self

    def make_output_dirs(self) -> None:
        if self.output_dir is not None:
            self.output_dir.mkdir(exist_ok=True)
        for name in DEFAULT_OUT_DIRS:
            d = getattr(self, name)
            if d:
                Path(d).mkdir(exist_ok=True)


def run_analysis(cfg: RNACliqueConfig, output_config: Optional[Path] = None):
    """rna_clique.py:180-217 without the argument parser: returns (SampleSimilarity, matrix DataFrame)."""
    from .rna_clique import rna_clique, write_matrix
    cfg.with_defaults()
    if not cfg.input_dirs or cfg.top_genes is None:
        raise YAMLConfigFileError("input_dirs and top_genes are required to run an analysis.")
    for d in cfg.input_dirs:
        if not (Path(d) / cfg.transcripts_name).is_file():
            raise FileNotFoundError(f"{Path(d) / cfg.transcripts_name} does not exist.")
    cfg.make_output_dirs()
    sim, pts = rna_clique(cfg.input_dirs, cfg.top_genes_dir, cfg.tables_dir, cfg.cache_dir, cfg.graph, None,
                          cfg.top_genes, cfg.transcripts_name, cfg.top_matches,
                          TranscriptID.parser_from_re(cfg.transcript_id_regex), cfg.evalue, cfg.keep_all, False,
                          jobs=cfg.jobs)
    cfg.path_to_sample = pts
    mat = sim.get_dissimilarity_df()
    write_matrix(mat, cfg.matrix)
    cfg.mark_finish()
    if output_config is not None:
        cfg.yaml_save(output_config)
    return sim, mat
