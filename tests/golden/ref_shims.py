"""Test-only shims that let the REFERENCE's own phase-2 / table-assembly / top-gene code be imported
in the build container (where multiset_key_dict, more_itertools, simple_blast and Biopython are
absent).  Used only by make_golden.py to generate the committed fixtures; never imported by the
product and never needed on the GPU box."""
import importlib.metadata
import sys
import types
from pathlib import Path

REF_SRC = Path("/root/reference/src")


class FrozenMultiset(tuple):
    def __new__(cls, it=()):
        return super().__new__(cls, sorted(it))


class MultisetKeyDict:
    """Minimal stand-in: a dict keyed by multisets (unordered pairs with multiplicity)."""

    def __class_getitem__(cls, item):
        return cls

    def __init__(self, it=()):
        self._d = {}
        if isinstance(it, MultisetKeyDict):
            self._d.update(it._d)
        else:
            items = it.items() if hasattr(it, "items") else it
            for k, v in items:
                self._d[FrozenMultiset(k)] = v

    def __getitem__(self, k):
        return self._d[FrozenMultiset(k)]

    def __setitem__(self, k, v):
        self._d[FrozenMultiset(k)] = v

    def items(self):
        return self._d.items()

    def multiset_iter(self):
        return iter(self._d.items())

    def __iter__(self):
        return iter(self._d.items())

    def key_elements(self):
        return {e for k in self._d for e in k}

    def __or__(self, other):
        r = MultisetKeyDict(self)
        r._d.update(other._d)
        return r

    def __len__(self):
        return len(self._d)


def install():
    if "multiset_key_dict" not in sys.modules:
        m = types.ModuleType("multiset_key_dict")
        m.MultisetKeyDict = MultisetKeyDict
        m.FrozenMultiset = FrozenMultiset
        sys.modules["multiset_key_dict"] = m
    if "more_itertools" not in sys.modules:
        m = types.ModuleType("more_itertools")

        def consume(it, n=None):
            for _ in it:
                pass
        m.consume = consume
        sys.modules["more_itertools"] = m
    if "simple_blast" not in sys.modules:
        m = types.ModuleType("simple_blast")
        b = types.ModuleType("simple_blast.blasting")

        class TabularBlastnSearch:  # never executed: gene_matches is replaced in the golden script
            def __init__(self, *a, **k):
                raise RuntimeError("BLAST+ is not available in this container")
        b.TabularBlastnSearch = TabularBlastnSearch
        m.blasting = b
        m.TabularBlastnSearch = TabularBlastnSearch
        m.BlastDBCache = object
        sys.modules["simple_blast"] = m
        sys.modules["simple_blast.blasting"] = b
    if "Bio" not in sys.modules:
        m = types.ModuleType("Bio")
        s = types.ModuleType("Bio.SeqIO")
        m.SeqIO = s
        m.SeqRecord = object
        sys.modules["Bio"] = m
        sys.modules["Bio.SeqIO"] = s
    real_version = importlib.metadata.version

    def version(name):
        if name == "rna_clique":
            return "0.3.0a3"
        return real_version(name)
    importlib.metadata.version = version
    if str(REF_SRC) not in sys.path:
        sys.path.insert(0, str(REF_SRC))
