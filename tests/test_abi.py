"""CPU: the C-ABI library loads and exports every symbol include/rcq.h declares (no compute calls)."""
import ctypes
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def header_symbols():
    text = (ROOT / "include" / "rcq.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rcq_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from rna_clique_b200 import _native
    assert _native.LIB_PATH.exists(), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(str(_native.LIB_PATH))
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} is declared in include/rcq.h but not exported"
    assert sorted(_native.SYMBOLS) == syms, "rna_clique_b200/_native.py SYMBOLS is out of date"


def test_host_only_entry_points():
    """The Karlin-Altschul helpers are plain host functions: usable without a GPU."""
    from rna_clique_b200 import _native
    from oracle import oracle as O
    L = _native.lib()
    for m, n, nseq in [(2050, 102_500_000, 50_000), (300, 1_000_000, 500), (15_000, 40_000, 20)]:
        assert L.rcq_length_adjustment(m, n, nseq) == O.lib().orc_length_adjustment(m, n, nseq)
        for s in (28, 198, 400, 2000):
            a, b = L.rcq_evalue(s, m, n, nseq), O.lib().orc_evalue(s, m, n, nseq)
            assert abs(a - b) <= 1e-12 * abs(b)      # exp(log(Kmn) - lambda S) vs Kmn exp(-lambda S)
    assert abs(L.rcq_bitscore(198) - 366.75) < 0.1


def test_no_device_fails_loudly():
    """There is no CPU fallback: creating a context without a CUDA device is an error."""
    import pytest
    import torch
    from rna_clique_b200 import _native
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_native.RcqError):
        _native.Context(0)


def test_product_does_not_import_oracle():
    for p in (ROOT / "rna_clique_b200").rglob("*.py"):
        src = p.read_text()
        assert "import oracle" not in src and "from oracle" not in src, p
