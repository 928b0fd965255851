"""Synthetic transcriptomes: a restatement of the reference's benchmark-input generator.

Reference: distance_sequence_simulator/simulation.py:344-388 (Simulation.simulate),
sim_config.yaml:1-39 (birth-death tree birth 1.0 / death 0.5; lengths 1950 + Binomial(1000, 0.1);
HKY85 with mutation_rate 0.01, dendropy defaults kappa=1 and equal base frequencies, i.e.
Jukes-Cantor; substitutions only; coverage U(0, 10000) drawn once per gene and shared by all
taxa; ids ``NODE_cov_{cov}_g{gene}_i1``; one ``T<k>/transcripts.fasta`` per taxon; seed 487).
dendropy and Biopython are not available, so the exact RNG streams differ: datasets are
"same distribution, our seed".  torch is used only as an array library (CPU or CUDA).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from pathlib import Path
from typing import Optional

import numpy as np
import torch


@dataclass
class Tree:
    """Rooted tree over the extant tips: parent[i] (root = -1), branch length to the parent."""
    parent: list
    length: list
    tips: list          # node index of tip k (taxon T<k>)

    def depth_order(self):
        order, children = [], {}
        for i, p in enumerate(self.parent):
            children.setdefault(p, []).append(i)
        stack = list(children.get(-1, []))
        while stack:
            n = stack.pop()
            order.append(n)
            stack.extend(children.get(n, []))
        return order

    def distance(self, a: int, b: int) -> float:
        """Path length between tips a and b."""
        def chain(n):
            out, d = {}, 0.0
            while n != -1:
                out[n] = d
                d += self.length[n]
                n = self.parent[n]
            return out
        ca, cb = chain(self.tips[a]), chain(self.tips[b])
        return min(ca[n] + cb[n] for n in ca if n in cb)


def birth_death_tree(n_tips: int, birth: float = 1.0, death: float = 0.5,
                     rng: Optional[np.random.Generator] = None) -> Tree:
    """Forward birth-death simulation stopped when n_tips lineages are extant
    (dendropy birth_death_tree(num_extant_tips=...), reference simulation_config.py tree_generator);
    extinct lineages are pruned because they leave no sequences."""
    rng = rng or np.random.default_rng(487)
    while True:
        parent, length, alive = [-1], [0.0], [0]
        ok = True
        while len(alive) < n_tips:
            k = len(alive)
            dt = rng.exponential(1.0 / (k * (birth + death)))
            for n in alive:
                length[n] += dt
            who = alive[rng.integers(k)]
            if rng.random() < birth / (birth + death):
                alive.remove(who)
                for _ in range(2):
                    parent.append(who); length.append(0.0); alive.append(len(parent) - 1)
            else:
                alive.remove(who)
                if not alive:
                    ok = False
                    break
        if not ok:
            continue
        dt = rng.exponential(1.0 / (len(alive) * (birth + death)))  # tips get some length too
        for n in alive:
            length[n] += dt
        break
    # prune to the tree induced by extant tips
    keep = set()
    for n in alive:
        while n != -1 and n not in keep:
            keep.add(n); n = parent[n]
    idx = {n: i for i, n in enumerate(sorted(keep))}
    new_parent = [idx.get(parent[n], -1) if parent[n] in keep else -1 for n in sorted(keep)]
    new_length = [length[n] for n in sorted(keep)]
    return Tree(new_parent, new_length, [idx[n] for n in alive])


def star_tree(n_tips: int, divergences) -> Tree:
    """Tips hanging off one root with given root-to-tip lengths (divergence sweeps,
    reference tree_gen.py:4-37)."""
    d = list(divergences)
    return Tree([-1] + [0] * n_tips, [0.0] + d, list(range(1, n_tips + 1)))


@dataclass
class Transcriptome:
    """One taxon's transcripts: concatenated ASCII letters + offsets + ids."""
    name: str
    ascii: torch.Tensor        # uint8, concatenated, no newlines
    offsets: torch.Tensor      # int64 [n+1] (CPU)
    coverage: np.ndarray       # float64 [n]
    gene: np.ndarray           # int32 [n]
    iso: np.ndarray            # int32 [n]

    def ids(self):
        return [f"NODE_cov_{float(c)!r}_g{int(g)}_i{int(i)}" for c, g, i in zip(self.coverage, self.gene, self.iso)]


_LETTERS = torch.tensor([65, 67, 71, 84], dtype=torch.uint8)  # A C G T


def simulate(taxa: int, count: int, seed: int = 487, mutation_rate: float = 0.01,
             tree: Optional[Tree] = None, device: str = "cpu", len_loc: int = 1950,
             len_n: int = 1000, len_p: float = 0.1, indel_rate: float = 0.0):
    """Returns (tree, [Transcriptome per taxon T0..T{taxa-1}]).

    indel_rate > 0 adds single-base deletions at that per-site per-unit-length rate (NOT in the
    reference's simulator, which is substitution-only; used by tests to exercise the gapped kernel)."""
    rng = np.random.default_rng(seed)
    if tree is None:
        tree = birth_death_tree(taxa, rng=rng)
    lens = len_loc + rng.binomial(len_n, len_p, size=count)
    cov = rng.uniform(0.0, 10000.0, size=count)
    offsets = np.zeros(count + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    total = int(offsets[-1])
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    root = torch.randint(0, 4, (total,), generator=g, device=device, dtype=torch.uint8)
    seqs = {-1: root}
    order = tree.depth_order()
    tipset = {n: k for k, n in enumerate(tree.tips)}
    out = [None] * taxa
    remaining_children = {}
    for p in tree.parent:
        remaining_children[p] = remaining_children.get(p, 0) + 1
    letters = _LETTERS.to(device)
    for n in order:
        par = seqs[tree.parent[n]]
        d = mutation_rate * tree.length[n]
        p_diff = 0.75 * (1.0 - math.exp(-4.0 * d / 3.0))
        u = torch.rand(total, generator=g, device=device)
        shift = torch.randint(1, 4, (total,), generator=g, device=device, dtype=torch.uint8)
        child = torch.where(u < p_diff, (par + shift) & 3, par)
        del u, shift
        seqs[n] = child
        remaining_children[tree.parent[n]] -= 1
        if remaining_children[tree.parent[n]] == 0:
            del seqs[tree.parent[n]]
        if n in tipset:
            k = tipset[n]
            codes, offs = child, offsets
            if indel_rate > 0:
                keep = torch.rand(total, generator=g, device=device) >= indel_rate * tree.length[n]
                keep_cpu = keep.cpu().numpy()
                csum = np.concatenate([[0], np.cumsum(keep_cpu)])
                offs = csum[offsets].astype(np.int64)
                codes = child[keep]
            out[k] = Transcriptome(f"T{k}", letters[codes.long()], torch.from_numpy(np.asarray(offs)),
                                   cov.copy(), np.arange(count, dtype=np.int32), np.ones(count, dtype=np.int32))
            if remaining_children.get(n, 0) == 0:
                del seqs[n]
    return tree, out


def write_fasta(t: Transcriptome, path: Path, width: int = 60) -> None:
    """Biopython-style FASTA (60 columns), as Simulation.simulate writes it (simulation.py:380-386)."""
    a = t.ascii.cpu().numpy()
    offs = t.offsets.numpy()
    ids = t.ids()
    with open(path, "wb") as f:
        for i, name in enumerate(ids):
            s = a[offs[i]:offs[i + 1]]
            f.write(b">" + name.encode() + b"\n")
            n = len(s)
            full = (n // width) * width
            if full:
                body = np.empty((n // width, width + 1), dtype=np.uint8)
                body[:, :width] = s[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(s[full:].tobytes() + b"\n")


def write_dataset(out_dir: Path, transcriptomes, transcripts_name: str = "transcripts.fasta"):
    """One directory per taxon holding transcripts.fasta (simulation.py:300-304). Returns dirs."""
    out_dir = Path(out_dir)
    dirs = []
    for t in transcriptomes:
        d = out_dir / t.name
        d.mkdir(parents=True, exist_ok=True)
        write_fasta(t, d / transcripts_name)
        dirs.append(d)
    return dirs
