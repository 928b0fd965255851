"""Generates the committed golden fixtures by running the REFERENCE's own code (from
/root/reference, with the shims of ref_shims.py) on synthetic inputs.  Run in the build
container only:  python tests/golden/make_golden.py

  phase2_*.json      reference build_graph + SampleSimilarity on synthetic gene-matches tables
                     -> valid (sample, gene) set, per-pair Fraction, fp64 dissimilarity matrix
  match_table_*.json reference HomologFinder.get_match_table driven by synthetic BLAST hit frames
  top_genes.json     reference TopGeneSelector on synthetic FASTA ids
"""
import json
import sys
from fractions import Fraction
from pathlib import Path

import numpy as np
import pandas as pd

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import ref_shims  # noqa: E402

ref_shims.install()
from rna_clique.build_graph import build_graph  # noqa: E402
from rna_clique.filtered_distance import SampleSimilarity, NoIdealComponentsError  # noqa: E402
from rna_clique.similarity_computer import ComparisonSimilarityComputer  # noqa: E402
from rna_clique import find_homologs as fh  # noqa: E402
from rna_clique.select_top_genes import TopGeneSelector  # noqa: E402
from rna_clique.transcripts import TranscriptID, default_gene_re  # noqa: E402


# ------------------------------------------------------------------ phase 2
def synth_tables(rng, n_samples, n_genes, p_drop=0.1, p_paralog=0.05, p_tie=0.3, p_wrong=0.03, shuffle=True):
    """Synthetic gene-matches tables: ortholog groups with drop-out, paralog fan-out, wrong edges
    and forward/backward tie rows (the duplicate rows the reference produces)."""
    names = [f"out/od1/S{chr(ord('a') + (7 * i) % 26)}{i}_top.fasta" for i in range(n_samples)]
    tables = []
    for i in range(n_samples):
        for j in range(i + 1, n_samples):
            rows = []
            for g in range(n_genes):
                if rng.random() < p_drop:
                    continue
                sg, qg = g, g
                if rng.random() < p_wrong:
                    qg = int(rng.integers(0, n_genes))
                length = int(rng.integers(300, 3000))
                gaps = int(rng.integers(0, 5)) if rng.random() < 0.2 else 0
                nident = length - gaps - int(rng.integers(0, 60))
                bits = int(1.85 * (2 * nident - length))
                row = dict(nident=nident, length=length, gaps=gaps, bitscore=float(bits), sgene=sg, qgene=qg,
                           reverse=False)
                rows.append(row)
                if rng.random() < p_tie:
                    rows.append(dict(row, reverse=True))
                if rng.random() < p_paralog:
                    rows.append(dict(row, sgene=n_genes + int(rng.integers(0, 5)), reverse=bool(rng.integers(0, 2))))
            if shuffle:
                rng.shuffle(rows)
            df = pd.DataFrame(rows, columns=["nident", "length", "gaps", "bitscore", "sgene", "qgene", "reverse"])
            df["ssample"] = names[i]
            df["qsample"] = names[j]
            tables.append(df)
    return names, tables


def run_phase2(tables, sample_count=None):
    graph = build_graph(tables)
    sim = SampleSimilarity(graph, ComparisonSimilarityComputer.mapping_from_dfs(iter(tables)), sample_count)
    out = {"n_nodes": graph.number_of_nodes(), "n_edges": graph.number_of_edges()}
    try:
        df = sim.get_dissimilarity_df()
        out["samples"] = list(df.index)
        out["matrix"] = [[float(x) for x in row] for row in df.to_numpy()]
        sims = sim.similarities
        out["fractions"] = [[list(k), int(v.numerator), int(v.denominator)] for k, v in sims.items() if len(set(k)) == 2]
        out["error"] = None
    except NoIdealComponentsError:
        out["error"] = "NoIdealComponentsError"
    out["sample_count"] = int(sim.sample_count)
    out["valid"] = sorted([str(s), int(g)] for s, g in sim.valid.itertuples(index=False))
    return out


def tables_to_json(tables):
    return [{c: (df[c].tolist() if c not in ("ssample", "qsample") else (df[c].iloc[0] if len(df) else None))
             for c in df.columns} for df in tables]


def make_phase2():
    cases = {
        "small": dict(seed=1, n_samples=3, n_genes=12),
        "tie_heavy": dict(seed=2, n_samples=4, n_genes=40, p_tie=0.9),
        "noisy": dict(seed=3, n_samples=5, n_genes=80, p_drop=0.05, p_paralog=0.05, p_wrong=0.03),
        "very_noisy": dict(seed=8, n_samples=5, n_genes=60, p_drop=0.25, p_paralog=0.2, p_wrong=0.15),
        "clean": dict(seed=4, n_samples=6, n_genes=30, p_drop=0.0, p_paralog=0.0, p_wrong=0.0),
        "no_ideal": dict(seed=5, n_samples=3, n_genes=6, p_drop=0.9),
        "wide": dict(seed=6, n_samples=9, n_genes=40, p_drop=0.005, p_paralog=0.005, p_wrong=0.003),
    }
    for name, kw in cases.items():
        rng = np.random.default_rng(kw.pop("seed"))
        names, tables = synth_tables(rng, **kw)
        tables = [t for t in tables if len(t)]
        res = run_phase2(tables)
        (HERE / f"phase2_{name}.json").write_text(json.dumps({"names": names, "tables": tables_to_json(tables), "expect": res}))
        print("phase2", name, res["error"], res["sample_count"], len(res["valid"]))
    # explicit sample_count larger than the samples present
    rng = np.random.default_rng(7)
    names, tables = synth_tables(rng, 3, 10)
    res = run_phase2(tables, sample_count=4)
    (HERE / "phase2_count_override.json").write_text(json.dumps({"names": names, "tables": tables_to_json(tables), "expect": res, "sample_count": 4}))


# ------------------------------------------------------------------ get_match_table
HIT_COLS = ["pident", "length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "evalue", "bitscore",
            "gaps", "nident", "sstrand", "qgene", "qiso", "sgene", "siso"]


def synth_hits(rng, n_q, n_s, n_hits, score_levels=6):
    """A blastn-like hit frame (after seqid parsing): many hits per query gene, bitscore ties."""
    rows = []
    for _ in range(n_hits):
        length = int(rng.integers(200, 2500))
        gaps = int(rng.integers(0, 4))
        mismatch = int(rng.integers(0, 40))
        nident = length - gaps - mismatch
        score = int(rng.integers(0, score_levels)) * 37 + 200
        rows.append(dict(pident=round(100 * nident / length, 3), length=length, mismatch=mismatch, gapopen=min(gaps, 2),
                         qstart=int(rng.integers(1, 50)), qend=int(rng.integers(2000, 2500)),
                         sstart=int(rng.integers(1, 50)), send=int(rng.integers(2000, 2500)), evalue=0.0,
                         bitscore=float(score), gaps=gaps, nident=nident, sstrand=["plus", "minus"][int(rng.integers(0, 2))],
                         qgene=int(rng.integers(0, n_q)), qiso=int(rng.integers(1, 3)),
                         sgene=int(rng.integers(0, n_s)), siso=int(rng.integers(1, 3))))
    return pd.DataFrame(rows, columns=HIT_COLS)


def ref_match_table(fwd_hits, bwd_hits, top_n, keep_all):
    hf = fh.HomologFinder(TranscriptID.parser_from_re(default_gene_re), top_n, 1e-99, keep_all)
    frames = {"f": fwd_hits, "b": bwd_hits}

    def gm(path1, path2):
        hits = frames[path1].copy()
        res = fh.highest_bitscores(hits, top_n, keep="all")
        return fh.shrink_df(res)
    hf.gm = gm
    return hf.get_match_table("f", "b")


def make_match_tables():
    cases = {
        "basic": dict(seed=11, n_q=12, n_s=12, n_hits=90, top_n=1, keep_all=True),
        "ties": dict(seed=12, n_q=8, n_s=8, n_hits=160, top_n=1, keep_all=True, score_levels=3),
        "first": dict(seed=13, n_q=8, n_s=8, n_hits=160, top_n=1, keep_all=False, score_levels=3),
        "sparse": dict(seed=14, n_q=30, n_s=30, n_hits=60, top_n=1, keep_all=True),
        "top2": dict(seed=15, n_q=8, n_s=8, n_hits=140, top_n=2, keep_all=True, score_levels=4),
    }
    for name, kw in cases.items():
        rng = np.random.default_rng(kw["seed"])
        f = synth_hits(rng, kw["n_q"], kw["n_s"], kw["n_hits"], kw.get("score_levels", 6))
        b = synth_hits(rng, kw["n_s"], kw["n_q"], kw["n_hits"], kw.get("score_levels", 6))
        # make the two directions share gene pairs (reciprocal hits), some with equal scores
        share = f.sample(frac=0.7, random_state=kw["seed"]).copy()
        share[["qgene", "sgene"]] = share[["sgene", "qgene"]].to_numpy()
        share[["qiso", "siso"]] = share[["siso", "qiso"]].to_numpy()
        bump = rng.integers(0, 3, size=len(share)) - 1
        share["bitscore"] = share["bitscore"] + 37.0 * bump
        b = pd.concat([b, share]).reset_index(drop=True)
        out = ref_match_table(f, b, kw["top_n"], kw["keep_all"])
        rec = {"top_n": kw["top_n"], "keep_all": kw["keep_all"],
               "fwd": {c: f[c].tolist() for c in HIT_COLS}, "bwd": {c: b[c].tolist() for c in HIT_COLS},
               "expect": {c: out[c].tolist() for c in out.columns}, "expect_index": [int(i) for i in out.index]}
        (HERE / f"match_table_{name}.json").write_text(json.dumps(rec))
        print("match_table", name, len(out))


# ------------------------------------------------------------------ top genes
class _Rec:
    def __init__(self, id_):
        self.id = id_


def make_top_genes():
    rng = np.random.default_rng(21)
    ids = []
    for g in range(40):
        for iso in range(1, int(rng.integers(1, 4)) + 1):
            cov = float(rng.integers(1, 12)) + (0.5 if rng.random() < 0.5 else 0.25)   # many coverage ties
            ids.append(f"NODE_{len(ids)}_length_{int(rng.integers(200, 900))}_cov_{cov!r}_g{g}_i{iso}")
    rng.shuffle(ids)
    out = {"ids": ids, "cases": []}
    for top in (1, 5, 17, 40, 100):
        sel = TopGeneSelector(lambda: iter([_Rec(i) for i in ids]), top)
        genes = sorted(int(g) for g in sel.get_top_genes())
        kept = [r.id for r in sel.get_top_gene_seqs()]
        out["cases"].append({"top": top, "genes": genes, "kept_ids": kept})
    (HERE / "top_genes.json").write_text(json.dumps(out))
    print("top_genes", [len(c["kept_ids"]) for c in out["cases"]])


if __name__ == "__main__":
    make_phase2()
    make_match_tables()
    make_top_genes()
