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


_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def _mutate(rng, codes, sub, indel):
    """Substitutions at rate `sub`, single-base deletions and insertions at rate `indel` each."""
    c = codes.copy()
    hit = rng.random(len(c)) < sub
    c[hit] = (c[hit] + rng.integers(1, 4, size=int(hit.sum()))) & 3
    if indel > 0:
        c = c[rng.random(len(c)) >= indel]
        ins = np.flatnonzero(rng.random(len(c)) < indel)
        c = np.insert(c, ins, rng.integers(0, 4, size=len(ins)))
    return c


def make_family_samples(seed=1, genes=60, taxa=2, sub=0.03, indel=0.003, length=(700, 1600), tails=0):
    """Samples on which the table assembly has real decisions to take (find_homologs.py:135-164, 215-302):

      * 1-3 isoforms per gene: the full transcript, one with an internal exon skipped, one truncated, in
        shuffled file order and with different sets per sample;
      * paralog families: a gene copied inside every sample with 1-5 % extra divergence under its own gene
        id (one query gene -> several subject genes), some copies EXACT (equal scores across genes) and some
        present in one sample only (best hits that are not reciprocal);
      * chimeric transcripts joining halves of two genes;
      * sparse, sample-specific gene ids (the dense gene index differs from the id, and an ortholog has a
        different id in every sample), isoform numbers that do not start at 1;
      * tails > 0: that many transcripts get a poly-A tail or a (CA)n stretch (low-complexity seeds).

    Returns [(ascii, offsets, gene, iso)] per taxon, the form make_samples returns."""
    rng = np.random.default_rng(seed)
    roots = [rng.integers(0, 4, size=int(rng.integers(*length))).astype(np.uint8) for _ in range(genes)]
    # family plan shared by all samples: (root gene, extra divergence, which samples carry it)
    plan = [(g, 0.0, None) for g in range(genes)]
    for g in rng.choice(genes, size=max(genes // 5, 1), replace=False):
        kind = rng.integers(0, 3)
        div = 0.0 if kind == 0 else float(rng.uniform(0.01, 0.05))
        only = None if kind < 2 else int(rng.integers(0, taxa))
        plan.append((int(g), div, only))
    para_seq = {}
    out = []
    for k in range(taxa):
        ids = rng.choice(np.arange(3, 400_000), size=len(plan), replace=False)     # sparse, per sample
        recs = []
        sample_full = {}
        for slot, (g, div, only) in enumerate(plan):
            if only is not None and only != k:
                continue
            base = roots[g]
            if slot >= genes:
                if slot not in para_seq:
                    para_seq[slot] = _mutate(rng, base, div, 0.0)
                base = para_seq[slot]
            if slot >= genes and div == 0.0:
                full = sample_full[g]           # exact duplicate inside the sample: equal scores across genes
            else:
                full = _mutate(rng, base, sub, indel)
            if slot < genes:
                sample_full[g] = full
            iso0 = int(rng.integers(1, 4))
            forms = [full]
            if rng.random() < 0.5 and len(full) > 500:
                a = int(rng.integers(150, len(full) // 2))
                b = a + int(rng.integers(60, 200))
                forms.append(np.concatenate([full[:a], full[b:]]))           # exon skipped
            if rng.random() < 0.4 and len(full) > 500:
                forms.append(full[int(rng.integers(0, 200)):len(full) - int(rng.integers(0, 200))])   # truncated
            if rng.random() < 0.5:
                forms = forms[::-1]
            for j, f in enumerate(forms):
                if rng.random() < 0.3:
                    f = (3 - f[::-1]).astype(np.uint8)                         # assembled on the other strand
                recs.append((int(ids[slot]), iso0 + j, f))
        for _ in range(max(genes // 15, 1)):                                    # chimeras
            a, b = rng.choice(genes, size=2, replace=False)
            fa, fb = _mutate(rng, roots[a], sub, 0.0), _mutate(rng, roots[b], sub, 0.0)
            recs.append((int(rng.integers(400_000, 500_000)), 1, np.concatenate([fa[:len(fa) // 2], fb[len(fb) // 2:]])))
        order = rng.permutation(len(recs))
        recs = [recs[i] for i in order]
        for t in range(min(tails, len(recs))):
            g_, i_, f = recs[t]
            tail = np.zeros(int(rng.integers(30, 70)), dtype=np.uint8) if t % 2 == 0 else \
                np.tile(np.array([1, 0], dtype=np.uint8), int(rng.integers(18, 40)))
            recs[t] = (g_, i_, np.concatenate([f, tail]))
        seq = np.concatenate([f for _, _, f in recs])
        offs = np.concatenate([[0], np.cumsum([len(f) for _, _, f in recs])]).astype(np.int64)
        out.append((_ACGT[seq], offs, np.array([r[0] for r in recs], dtype=np.int32),
                    np.array([r[1] for r in recs], dtype=np.int32)))
    return out
