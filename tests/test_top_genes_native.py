"""rcq_top_gene_mask (host code of the native library) against the numpy statement of
select_top_genes.py:108-127 and against the reference-generated fixture."""
import json
from pathlib import Path

import numpy as np

from rna_clique_b200.select_top_genes import top_gene_mask, top_gene_mask_numpy


def test_native_equals_numpy_on_random_cases():
    rng = np.random.default_rng(0)
    for trial in range(300):
        n = int(rng.integers(1, 400))
        gene = rng.integers(0, max(1, n // 2) + (10 ** 9 if trial % 7 == 0 else 0), size=n).astype(np.int32)
        cov = rng.integers(0, 6, size=n).astype(np.float64) if trial % 2 else rng.uniform(0, 100, size=n)
        top = int(rng.integers(-1, n + 2))
        assert np.array_equal(top_gene_mask(cov, gene, top), top_gene_mask_numpy(cov, gene, top)), trial


def test_ties_go_to_the_larger_gene_id():
    cov = np.array([5.0, 5.0, 5.0, 1.0])
    gene = np.array([3, 9, 4, 100], dtype=np.int32)
    assert top_gene_mask(cov, gene, 2).tolist() == [1, 2]        # genes 9 and 4 beat gene 3
    assert top_gene_mask(cov, gene, 0).tolist() == []
    assert top_gene_mask(cov, gene, 10).tolist() == [0, 1, 2, 3]
    assert top_gene_mask(np.zeros(0), np.zeros(0, dtype=np.int32), 3).tolist() == []


def test_all_isoforms_of_a_selected_gene_are_kept():
    cov = np.array([1.0, 9.0, 2.0, 3.0])
    gene = np.array([7, 7, 8, 8], dtype=np.int32)
    assert top_gene_mask(cov, gene, 1).tolist() == [0, 1]
