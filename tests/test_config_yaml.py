"""The analysis config YAML (reference config.py:128-169): same keys, unknown keys rejected, None hidden
on save, defaults derived from output_dir with the reference's names (config.py:1472-1483)."""
import re
from pathlib import Path

import pytest
import yaml

from rna_clique_b200.config import RNACliqueConfig, YAMLConfigFileError


def test_round_trip_and_defaults(tmp_path):
    cfg = RNACliqueConfig(title="t", input_dirs=[tmp_path / "a", tmp_path / "b"], output_dir=tmp_path / "out",
                          top_genes=50000, transcript_id_regex=re.compile(r"cov_([0-9.]+)_g([0-9]+)_i([0-9]+)"))
    p = tmp_path / "config.yaml"
    cfg.yaml_save(p)
    raw = yaml.safe_load(p.read_text())
    assert "graph" not in raw and raw["top_genes"] == 50000 and raw["input_dirs"] == [str(tmp_path / "a"), str(tmp_path / "b")]
    assert raw["transcript_id_regex"] == r"cov_([0-9.]+)_g([0-9]+)_i([0-9]+)" and raw["evalue"] == 1e-99 and raw["keep_all"] is True
    back = RNACliqueConfig.yaml_load(p)
    assert back.input_dirs == cfg.input_dirs and back.output_dir == tmp_path / "out" and back.top_matches == 1
    assert back.transcript_id_regex.pattern == cfg.transcript_id_regex.pattern
    back.with_defaults()
    assert (back.top_genes_dir, back.tables_dir, back.cache_dir) == (tmp_path / "out" / "od1", tmp_path / "out" / "od2", tmp_path / "out" / "db_cache")
    assert back.graph == tmp_path / "out" / "graph.pkl" and back.matrix.stem == "distance_matrix"
    back.make_output_dirs()
    assert (tmp_path / "out" / "od2").is_dir()


def test_unknown_key_and_non_mapping_are_rejected(tmp_path):
    p = tmp_path / "bad.yaml"
    p.write_text("top_genes: 10\nnot_a_setting: 3\n")
    with pytest.raises(YAMLConfigFileError, match="Unknown configuration setting not_a_setting"):
        RNACliqueConfig.yaml_load(p)
    p.write_text("- 1\n- 2\n")
    with pytest.raises(YAMLConfigFileError, match="does not contain a mapping"):
        RNACliqueConfig.yaml_load(p)
    p.write_text("transcript_id_regex: '(['\n")
    with pytest.raises(YAMLConfigFileError, match="Failed to load"):
        RNACliqueConfig.yaml_load(p)
