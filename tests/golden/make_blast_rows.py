"""Extracts the example gene-matches rows of the reference's docs/formats.md (real blastn megablast
output on rnaSPAdes assemblies, lines 260-273) into tests/golden/blast_rows_formats_md.json.

These are the only rows of real BLAST+ output the reference holds.  They pin, independently of this
repository's own restatement: the raw-score formula of the +1/-2, linear-gap-2.5 scoring system, the
Karlin-Altschul constants lambda = 1.28 / K = 0.46 behind `bitscore`, BLAST's integer printing of bit
scores above 99.9 (truncation), pident to three decimals and the printed e-value (three significant
digits, 0.0 below 1e-180).  Run in the build container only (reads /root/reference)."""
import json
import re
from pathlib import Path

SRC = Path("/root/reference/docs/formats.md")
OUT = Path(__file__).resolve().parent / "blast_rows_formats_md.json"


def cell(x: str):
    x = x.strip()
    if x.startswith("`"):
        return x.strip("`")
    if x in ("True", "False"):
        return x == "True"
    x = x.strip("$").strip()
    if x == "":
        return 0                      # DataFrame.to_markdown leaves zeros blank in this table
    m = re.fullmatch(r"([0-9.]+) \\times 10\^\{(-?[0-9]+)\}", x)
    if m:
        return float(f"{m.group(1)}e{m.group(2)}")
    return float(x) if "." in x else int(x)


def main():
    lines = SRC.read_text().splitlines()
    head = next(i for i, l in enumerate(lines) if l.startswith("| pident"))
    cols = [c.strip() for c in lines[head].strip("|").split("|")]
    rows = []
    for l in lines[head + 2:]:
        if not l.startswith("|"):
            break
        rows.append(dict(zip(cols, (cell(c) for c in l.strip().strip("|").split("|")))))
    OUT.write_text(json.dumps({"source": "docs/formats.md:260-273 (actapia/rna_clique @ a55e40a)", "rows": rows}, indent=1) + "\n")
    print(len(rows), "rows ->", OUT)


if __name__ == "__main__":
    main()
