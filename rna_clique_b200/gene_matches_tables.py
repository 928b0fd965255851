"""Gene-matches table files (reference gene_matches_tables.py:8-82): `.pkl` (pandas pickle) or
`.h5` (key "gene_matches", format "table"; needs PyTables, which this image lacks — `.h5` then
raises pandas' own ImportError, exactly as the reference would)."""
from __future__ import annotations

import itertools
from pathlib import Path
from typing import Iterable, Iterator

import pandas as pd


def have_pytables() -> bool:
    try:
        import tables  # noqa: F401
        return True
    except Exception:  # noqa: BLE001
        return False


def default_extension() -> str:
    """The reference always writes `.h5` (find_all_pairs.py:212-213); without PyTables the drop-in
    writes the reference's other supported format."""
    return "h5" if have_pytables() else "pkl"


def read_table(path: Path, head: int = None, head_unsupported: bool = True) -> pd.DataFrame:
    path = Path(path)
    if path.suffix == ".pkl":
        res = pd.read_pickle(path)
    elif path.suffix == ".h5":
        res = pd.read_hdf(path, stop=head)
    else:
        raise ValueError(f"Could not determine file type for extension {path.suffix}.")
    if head and head_unsupported:
        res = res.head(head)
    return res


def write_table(df: pd.DataFrame, path: Path):
    path = Path(path)
    if path.suffix == ".pkl":
        df.to_pickle(path)
    elif path.suffix == ".h5":
        df.to_hdf(path, key="gene_matches", format="table")
    else:
        raise ValueError(f"Could not determine file type for extension {path.suffix}.")


def multi_glob(path: Path, globs: Iterable[str]) -> Iterator[Path]:
    return itertools.chain.from_iterable(Path(path).glob(g) for g in globs)


def get_table_files(path: Path) -> Iterator[Path]:
    return multi_glob(path, ["*.h5", "*.pkl"])
