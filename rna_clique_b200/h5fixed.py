"""matrix.h5 without PyTables: the distance matrix in pandas' "fixed" HDF5 layout (reference
rna_clique.py:175-177 `mat.to_hdf(output_matrix, key="matrix")`, docs/formats.md:346-352).

write_frame -> rcq_matrix_h5_write (csrc/host_h5.cu): superblock 0, old-style groups, contiguous arrays, the
PyTables and pandas attributes.  read_frame is a small reader of exactly that subset of HDF5 (superblock 0/1,
version-1 object headers, symbol-table groups with one B-tree leaf level per group, contiguous layout,
fixed-length strings, 64-bit floats and integers), so that the package can read its own matrix back where
PyTables is missing; with PyTables installed `pd.read_hdf(path, key)` reads the same file."""
from __future__ import annotations

import ctypes as C
import struct
from pathlib import Path

import numpy as np
import pandas as pd

from . import _native

UNDEF = 0xFFFFFFFFFFFFFFFF


def write_frame(path, key: str, df: pd.DataFrame) -> None:
    """A float DataFrame with string labels as an HDF5 file holding one pandas fixed-format frame under `key`."""
    values = np.ascontiguousarray(df.to_numpy(dtype=np.float64))
    rows = [str(x).encode("utf-8") for x in df.index]
    cols = [str(x).encode("utf-8") for x in df.columns]
    ra = (C.c_char_p * len(rows))(*rows)
    ca = (C.c_char_p * len(cols))(*cols)
    _native.check(_native.lib().rcq_matrix_h5_write(str(path).encode(), key.strip("/").encode(), values.shape[0], values.shape[1],
                                                   C.cast(ra, C.c_void_p), C.cast(ca, C.c_void_p), values.ctypes.data))


class _File:
    def __init__(self, data: bytes):
        self.d = data
        if data[:8] != b"\x89HDF\r\n\x1a\n":
            raise ValueError("not an HDF5 file")
        ver = data[8]
        if ver not in (0, 1) or data[13] != 8 or data[14] != 8:
            raise ValueError("unsupported superblock (need version 0/1 with 8-byte offsets and lengths)")
        at = 24 + (4 if ver == 1 else 0)
        self.base, _, self.eof, _ = struct.unpack_from("<4Q", data, at)
        entry = at + 32
        self.root = self._entry(entry)

    def _entry(self, at):
        name_off, header, cache = struct.unpack_from("<QQI", self.d, at)
        btree, heap = struct.unpack_from("<QQ", self.d, at + 24)
        return dict(name_off=name_off, header=header, cache=cache, btree=btree, heap=heap)

    def messages(self, addr):
        """[(type, bytes)] of a version-1 object header (continuation blocks followed)."""
        d = self.d
        if d[addr] != 1:
            raise ValueError("unsupported object header version")
        n_msgs, = struct.unpack_from("<H", d, addr + 2)
        size, = struct.unpack_from("<I", d, addr + 8)
        blocks = [(addr + 16, size)]
        out = []
        while blocks and len(out) < n_msgs:
            at, left = blocks.pop(0)
            end = at + left
            while at + 8 <= end and len(out) < n_msgs:
                mtype, msize = struct.unpack_from("<HH", d, at)
                body = d[at + 8:at + 8 + msize]
                if mtype == 0x0010:
                    blocks.append(struct.unpack_from("<QQ", body, 0))
                out.append((mtype, body))
                at += 8 + msize
        return out

    def group_children(self, header_addr):
        """{name: symbol table entry} of an old-style group."""
        stab = next(b for t, b in self.messages(header_addr) if t == 0x0011)
        btree, heap = struct.unpack_from("<QQ", stab, 0)
        if self.d[heap:heap + 4] != b"HEAP":
            raise ValueError("bad local heap")
        heap_data, = struct.unpack_from("<Q", self.d, heap + 24)
        out = {}

        def walk(node):
            d = self.d
            if d[node:node + 4] == b"SNOD":
                n, = struct.unpack_from("<H", d, node + 6)
                for i in range(n):
                    e = self._entry(node + 8 + 40 * i)
                    s = heap_data + e["name_off"]
                    out[d[s:d.index(b"\0", s)].decode()] = e
                return
            if d[node:node + 4] != b"TREE" or d[node + 4] != 0:
                raise ValueError("bad group B-tree node")
            used, = struct.unpack_from("<H", d, node + 6)
            for i in range(used):
                child, = struct.unpack_from("<Q", d, node + 24 + 8 + 16 * i)
                walk(child)

        walk(btree)
        return out

    @staticmethod
    def _dtype(body):
        cls, ver = body[0] & 15, body[0] >> 4
        size, = struct.unpack_from("<I", body, 4)
        if ver != 1:
            raise ValueError("unsupported datatype version")
        if cls == 3:
            return np.dtype(f"S{size}"), bool(body[1] >> 4)
        if cls == 1 and size == 8:
            return np.dtype("<f8"), False
        if cls == 0:
            return np.dtype(("<i" if body[1] & 8 else "<u") + str(size)), False
        raise ValueError(f"unsupported datatype class {cls}")

    @staticmethod
    def _dims(body):
        if body[0] != 1:
            raise ValueError("unsupported dataspace version")
        return struct.unpack_from(f"<{body[1]}Q", body, 8)

    def attributes(self, header_addr):
        out = {}
        for t, b in self.messages(header_addr):
            if t != 0x000C:
                continue
            if b[0] != 1:
                raise ValueError("unsupported attribute version")
            nl, tl, sl = struct.unpack_from("<HHH", b, 2)
            p8 = lambda x: (x + 7) & ~7
            at = 8
            name = b[at:at + nl - 1].decode()
            at += p8(nl)
            dt, utf8 = self._dtype(b[at:at + tl])
            at += p8(tl)
            dims = self._dims(b[at:at + sl])
            at += p8(sl)
            n = int(np.prod(dims)) if dims else 1
            v = np.frombuffer(b, dtype=dt, count=n, offset=at)
            if not dims:
                v = v[0]
                if dt.kind == "S":
                    v = bytes(v).rstrip(b"\0")
                    v = v.decode("utf-8") if utf8 else v
                elif dt.kind in "iu":
                    v = int(v)
            out[name] = v
        return out

    def dataset(self, header_addr):
        msgs = dict((t, b) for t, b in self.messages(header_addr) if t in (1, 3, 8))
        dt, _ = self._dtype(msgs[3])
        dims = self._dims(msgs[1])
        lay = msgs[8]
        if lay[0] != 3 or lay[1] != 1:
            raise ValueError("only contiguous layout (version 3) is supported")
        addr, size = struct.unpack_from("<QQ", lay, 2)
        n = int(np.prod(dims)) if dims else 1
        if size != n * dt.itemsize or addr + size > len(self.d):
            raise ValueError("dataset does not fit the file")
        return np.frombuffer(self.d, dtype=dt, count=n, offset=addr).reshape(dims)


def read_frame(path, key: str = "matrix") -> pd.DataFrame:
    """The DataFrame stored by write_frame (or by pandas' own fixed format when it used contiguous arrays)."""
    f = _File(Path(path).read_bytes())
    node = f.root
    for part in key.strip("/").split("/"):
        kids = f.group_children(node["header"])
        if part not in kids:
            raise KeyError(f"No object named {key} in the file")
        node = kids[part]
    attrs = f.attributes(node["header"])
    if attrs.get("pandas_type") != "frame" or int(attrs.get("nblocks", 0)) != 1:
        raise ValueError("not a single-block pandas fixed-format frame")
    kids = f.group_children(node["header"])

    def index(name):
        a = f.attributes(kids[name]["header"])
        if a.get("kind") != "string":
            raise ValueError(f"{name}: only string labels are supported")
        return [x.rstrip(b"\0").decode(attrs.get("encoding", "UTF-8")) for x in f.dataset(kids[name]["header"])]

    cols, rows, items = index("axis0"), index("axis1"), index("block0_items")
    v = f.dataset(kids["block0_values"]["header"])
    if f.attributes(kids["block0_values"]["header"]).get("transposed"):
        v = v.T
    return pd.DataFrame(np.array(v.T), index=rows, columns=items)[cols]
