#!/usr/bin/env python
"""BLAST+ parity harness: runs wherever `blastn` is on PATH.

For a simulated sample pair (or any two FASTA files given with --fasta A B) it runs what the reference runs
(find_homologs.py:124 through simple_blast: `blastn` default task megablast, both strands, e-value cut, tabular
columns + gaps nident sstrand) in both directions, reduces the hits with the reference's own rules
(highest_bitscores :135-164, get_match_table :215-302, restated in oracle/rcq_oracle.c and pinned to the
reference by tests/golden/match_table_*.json), and compares with

  --against oracle   the CPU restatement (default; needs no GPU)
  --against gpu      rcq_search_pair through the C ABI

It prints the best-hit edge agreement (north star: >= 99.5 %), the per-column agreement of the rows both
sides have, and the absolute difference of the pair's distance (north star: <= 1e-3).  Exit code 0 when both
gates hold, 1 otherwise, 77 when blastn is absent (tests/test_blast_parity.py turns 77 into a skip).

  python tools/blast_parity.py [--count 300] [--mutation 0.03] [--indel 0.003] [--families] [--against gpu]"""
import argparse
import shutil
import subprocess
import sys
import tempfile
from fractions import Fraction
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

OUTFMT = "6 qseqid sseqid pident length mismatch gapopen qstart qend sstart send evalue bitscore gaps nident sstrand"


def write_fasta(path, raw):
    a, offs, gene, iso = raw
    with open(path, "wb") as f:
        for t in range(len(gene)):
            f.write(f">NODE_cov_1.0_g{int(gene[t])}_i{int(iso[t])}\n".encode())
            f.write(a[offs[t]:offs[t + 1]].tobytes() + b"\n")


def read_fasta(path):
    from rna_clique_b200.fasta import read_fasta as rf
    from rna_clique_b200.transcripts import default_parser, parse_fasta_ids
    fa = rf(Path(path))
    _, gene, iso = parse_fasta_ids(fa, default_parser)
    return np.asarray(fa.ascii), np.asarray(fa.offsets), gene, iso


def blastn_rows(query_fa, subject_fa, evalue, threads):
    """One blastn direction -> int32 rows in the oracle's 17-column layout (score from the integer columns)."""
    cmd = ["blastn", "-query", str(query_fa), "-subject", str(subject_fa), "-evalue", repr(evalue),
           "-outfmt", OUTFMT]
    if threads > 1:
        cmd += ["-num_threads", str(threads)]
    txt = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    rows = []
    import re
    pat = re.compile(r"_g([0-9]+)_i([0-9]+)")
    for line in txt.splitlines():
        f = line.split("\t")
        qg, qi = map(int, pat.search(f[0]).groups())
        sg, si = map(int, pat.search(f[1]).groups())
        length, mism, gapopen, qs, qe, ss, se = (int(x) for x in f[3:10])
        gaps, nident = int(f[12]), int(f[13])
        score = (2 * nident - 4 * mism - 5 * gaps) >> 1
        rows.append([length, mism, gapopen, qs, qe, ss, se, score, gaps, nident, qg, qi, sg, si, 0, int(f[14] == "minus"), 0])
    return np.asarray(rows, dtype=np.int32).reshape(-1, 17)


def distance(rows):
    return float(1 - Fraction(int(rows[:, 9].sum()), int((rows[:, 0] - rows[:, 8]).sum()))) if len(rows) else float("nan")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--count", type=int, default=300)
    ap.add_argument("--mutation", type=float, default=0.03)
    ap.add_argument("--indel", type=float, default=0.003)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--families", action="store_true", help="multi-isoform / paralog inputs of the selection tests")
    ap.add_argument("--fasta", nargs=2, metavar=("A", "B"))
    ap.add_argument("--evalue", type=float, default=1e-99)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--against", default="oracle", choices=["oracle", "gpu"])
    a = ap.parse_args()
    if shutil.which("blastn") is None:
        print("blastn is not on PATH: BLAST+ parity cannot be measured on this machine")
        return 77
    from helpers import make_family_samples, make_samples
    from oracle import oracle as O
    if a.fasta:
        raw = [read_fasta(p) for p in a.fasta]
    elif a.families:
        raw = make_family_samples(seed=a.seed, genes=a.count)
    else:
        raw = make_samples(taxa=2, count=a.count, seed=a.seed, mutation_rate=a.mutation, indel_rate=a.indel)[1]
    tmp = Path(tempfile.mkdtemp(prefix="rcq_blast_"))
    try:
        fa = [tmp / "s1.fasta", tmp / "s2.fasta"]
        for p, r in zip(fa, raw):
            write_fasta(p, r)
        # forward: query = sample 2, subject = sample 1; backward: the other way round (find_homologs.py:229-239)
        fwd = blastn_rows(fa[1], fa[0], a.evalue, a.threads)
        bwd = blastn_rows(fa[0], fa[1], a.evalue, a.threads)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    want = O.match_table_from_rows(fwd, bwd).rows(2)          # the reference's reduction of real BLAST hits
    if a.against == "oracle":
        got = O.search_pair(O.Sample(*raw[0]), O.Sample(*raw[1]), a.evalue).rows(2)
    else:
        from rna_clique_b200 import _native
        ctx = _native.Context(0)
        t = _native.search_pair(ctx, _native.Sample(ctx, *raw[0]), _native.Sample(ctx, *raw[1]), a.evalue)
        c = t.download()
        got = np.stack([c[k] for k in _native.TABLE_COLS] + [c["minus"].astype(np.int32), c["reverse"].astype(np.int32)], axis=1)
    ew = {(int(r[10]), int(r[12])) for r in want}
    eg = {(int(r[10]), int(r[12])) for r in got}
    agree = len(ew & eg) / max(len(ew | eg), 1)
    print(f"blastn rows: {len(fwd)} forward, {len(bwd)} backward -> table {len(want)} rows, {len(ew)} best-hit edges")
    print(f"{a.against} table {len(got)} rows, {len(eg)} edges; edge agreement {100 * agree:.3f} %  "
          f"(only BLAST {len(ew - eg)}, only {a.against} {len(eg - ew)})")
    # rows of edges both sides have, best row per edge: column agreement
    def best(rows):
        d = {}
        for r in rows.tolist():
            k = (r[10], r[12])
            if k not in d or r[7] > d[k][7]:
                d[k] = r
        return d
    bw, bg = best(want), best(got)
    names = ["length", "mismatch", "gapopen", "qstart", "qend", "sstart", "send", "score", "gaps", "nident"]
    common = sorted(ew & eg)
    for ci, n in enumerate(names):
        same = sum(bw[k][ci] == bg[k][ci] for k in common)
        print(f"  {n:9s} identical on {same}/{len(common)} common edges")
    dw, dg = distance(want), distance(got)
    print(f"distance: BLAST {dw:.9f}  {a.against} {dg:.9f}  |diff| {abs(dw - dg):.2e}")
    ok = agree >= 0.995 and abs(dw - dg) <= 1e-3
    print("PASS" if ok else "FAIL", "(gates: edge agreement >= 99.5 %, |distance diff| <= 1e-3)")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
