"""Shared helpers for the parity tests (oracle side + synthetic inputs)."""
import numpy as np

from rna_clique_b200.simulate import simulate, star_tree


def make_samples(taxa=3, count=200, seed=1, mutation_rate=0.01, indel_rate=0.0, tree=None, **kw):
    tree, ts = simulate(taxa, count, seed=seed, mutation_rate=mutation_rate, indel_rate=indel_rate, tree=tree, **kw)
    return tree, [(t.ascii.numpy(), t.offsets.numpy(), t.gene, t.iso) for t in ts]


def np_pack(ascii_, offsets):
    """Reference 2-bit codes / mask per base, per transcript (plain numpy)."""
    lut = np.full(256, 4, dtype=np.uint8)
    for ch, v in (("A", 0), ("C", 1), ("G", 2), ("T", 3), ("U", 3)):
        lut[ord(ch)] = v
        lut[ord(ch.lower())] = v
    return lut[ascii_]


def unpack_arena(words, off, length):
    """Bases [off, off+length) of a packed 2-bit arena as uint8 codes."""
    idx = np.arange(off, off + length)
    return ((words[idx >> 4] >> ((idx & 15) * 2).astype(np.uint32)) & 3).astype(np.uint8)


def unpack_mask(words, off, length):
    idx = np.arange(off, off + length)
    return ((words[idx >> 5] >> (idx & 31).astype(np.uint32)) & 1).astype(np.uint8)
