"""matrix.h5 (reference rna_clique.py:175-177, docs/formats.md:346-352) written by the library's own HDF5
writer: structure checked byte by byte against the HDF5 file-format specification's field layout, read back by
the package's reader, and by pandas itself wherever PyTables is importable."""
import struct

import numpy as np
import pandas as pd
import pytest

from rna_clique_b200 import h5fixed


def frame(n=6, seed=0):
    labels = [f"rnac_out/od1/SRR{2321383 + i}_top.fasta" for i in range(n)]
    m = np.random.default_rng(seed).random((n, n)) * 0.01
    m = (m + m.T) / 2
    np.fill_diagonal(m, 0.0)
    return pd.DataFrame(m, index=labels, columns=labels)


def test_round_trip_is_bit_exact(tmp_path):
    for n in (1, 2, 16, 200):
        df = frame(n, seed=n)
        p = tmp_path / f"m{n}.h5"
        h5fixed.write_frame(p, "matrix", df)
        back = h5fixed.read_frame(p, "matrix")
        assert list(back.index) == list(df.index) and list(back.columns) == list(df.columns)
        assert back.to_numpy().tobytes() == df.to_numpy().tobytes()          # fp64 bit for bit
    rect = pd.DataFrame(np.arange(6.0).reshape(2, 3), index=["r0", "r1"], columns=["x", "yy", "zzz é"])
    h5fixed.write_frame(tmp_path / "r.h5", "matrix", rect)
    assert h5fixed.read_frame(tmp_path / "r.h5").equals(rect)
    with pytest.raises(KeyError):
        h5fixed.read_frame(tmp_path / "r.h5", "other")


def test_file_structure_follows_the_hdf5_specification(tmp_path):
    df = frame(5)
    p = tmp_path / "m.h5"
    h5fixed.write_frame(p, "matrix", df)
    d = p.read_bytes()
    # superblock version 0: signature, versions, 8-byte offsets/lengths, group K values, addresses
    assert d[:8] == b"\x89HDF\r\n\x1a\n" and d[8:13] == bytes(5) and d[13] == 8 and d[14] == 8
    assert struct.unpack_from("<HHI", d, 16) == (4, 16, 0)
    base, freespace, eof, driver = struct.unpack_from("<4Q", d, 24)
    assert (base, freespace, eof, driver) == (0, h5fixed.UNDEF, len(d), h5fixed.UNDEF)
    f = h5fixed._File(d)
    assert f.root["cache"] == 1 and d[f.root["btree"]:f.root["btree"] + 4] == b"TREE" and d[f.root["heap"]:f.root["heap"] + 4] == b"HEAP"
    # every object header: version 1, message sizes multiples of 8 that add up to the header size
    def check_header(addr):
        assert d[addr] == 1 and addr % 8 == 0
        n, = struct.unpack_from("<H", d, addr + 2)
        size, = struct.unpack_from("<I", d, addr + 8)
        at, seen = addr + 16, 0
        while at < addr + 16 + size:
            t, s = struct.unpack_from("<HH", d, at)
            assert s % 8 == 0
            at += 8 + s
            seen += 1
        assert at == addr + 16 + size and seen == n
    check_header(f.root["header"])
    top = f.group_children(f.root["header"])
    assert list(top) == ["matrix"] and top["matrix"]["cache"] == 1
    check_header(top["matrix"]["header"])
    kids = f.group_children(top["matrix"]["header"])
    assert list(kids) == ["axis0", "axis1", "block0_items", "block0_values"]          # symbol table nodes are sorted by name
    for e in kids.values():
        check_header(e["header"])
        assert e["cache"] == 0
    # the attributes pandas' fixed-format reader and PyTables look at
    assert f.attributes(f.root["header"]) == {"CLASS": "GROUP", "VERSION": "1.0", "PYTABLES_FORMAT_VERSION": "2.1"}
    ga = f.attributes(top["matrix"]["header"])
    assert ga == {"CLASS": "GROUP", "VERSION": "1.0", "pandas_type": "frame", "pandas_version": "0.15.2", "encoding": "UTF-8",
                  "errors": "strict", "ndim": 2, "nblocks": 1, "axis0_variety": "regular", "axis1_variety": "regular",
                  "block0_items_variety": "regular"}
    for name in ("axis0", "axis1", "block0_items"):
        a = f.attributes(kids[name]["header"])
        assert a == {"CLASS": "ARRAY", "VERSION": "2.4", "FLAVOR": "numpy", "kind": "string", "name": b"N."}
        import pickle
        assert pickle.loads(a["name"]) is None                                       # how PyTables stores None
        arr = f.dataset(kids[name]["header"])
        assert arr.dtype.kind == "S" and arr.shape == (5,)
    v = f.dataset(kids["block0_values"]["header"])
    assert v.dtype == np.dtype("<f8") and v.shape == (5, 5) and np.array_equal(v.T, df.to_numpy())
    # B-tree node and symbol-table node sizes for K = 16 / 4; the local heap holds the names, 8-aligned
    bt = top["matrix"]["btree"]
    assert struct.unpack_from("<BBH", d, bt + 4) == (0, 0, 1) and struct.unpack_from("<QQ", d, bt + 8) == (h5fixed.UNDEF, h5fixed.UNDEF)
    heap = top["matrix"]["heap"]
    hsize, hfree, hdata = struct.unpack_from("<QQQ", d, heap + 8)
    assert hsize % 8 == 0 and hfree == 1 and d[hdata:hdata + 8] == bytes(8)
    assert d[hdata + 8:hdata + 14] == b"axis0\0"


def test_rna_clique_write_matrix_h5(tmp_path):
    from rna_clique_b200.rna_clique import write_matrix
    df = frame(4)
    out = write_matrix(df, tmp_path / "distance_matrix.h5")
    assert out.exists() and h5fixed.read_frame(out).equals(df)


def test_pandas_reads_the_file_when_pytables_is_installed(tmp_path):
    pytest.importorskip("tables")
    df = frame(7)
    p = tmp_path / "m.h5"
    h5fixed.write_frame(p, "matrix", df)
    back = pd.read_hdf(p, "matrix")
    assert back.equals(df)
    # and the other way round: pandas' own file, read by the package's reader
    q = tmp_path / "pandas.h5"
    df.to_hdf(q, key="matrix")
    assert h5fixed.read_frame(q).equals(df)
