#!/usr/bin/env python
"""Dynamic instruction counts per SOURCE LINE (or per opcode) of one kernel: joins the per-instruction
'Instructions Executed' column of an ncu report (--page source --csv, captured with --import-source on / -lineinfo)
with the line table nvdisasm prints for the same cubin.  The two listings are matched by instruction order.

    cuobjdump -xelf all rna_clique_b200/librnaclique_b200.so          # -> search.sm_100a.cubin ...
    python tools/sass_lines.py x.ncu-rep search.sm_100a.cubin k_gapped_reg --mangled k_gapped_regILi8ELi6E [--per N]

--per N divides the counts by N (e.g. the number of warp-waves) so that the figures read 'instructions per wave'.
"""
import argparse
import collections
import csv
import io
import re
import subprocess


def ncu_counts(rep, kernel_regex, skip):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kernel_regex}",
                          "--launch-skip", str(skip), "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    col = {n: i for i, n in enumerate(rows[hdr])}
    res = []
    for r in rows[hdr + 1:]:
        if r and r[0] in ("Kernel Name", "Address"):          # the next launch of the report
            break
        if len(r) <= col["Instructions Executed"]:
            continue
        res.append((r[col["Source"]].strip(), int(r[col["Instructions Executed"]]), int(r[col["# Samples"]] or 0)))
    return res


def disasm_lines(cubin, kernel_substr):
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    res, inside, cur = [], False, ("?", 0)
    for line in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", line)
        if m:
            inside = kernel_substr in m.group(1)
            continue
        if re.match(r"\s*\.section", line):
            inside = False
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if m:
            res.append((cur, m.group(1).strip()))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("cubin")
    ap.add_argument("kernel", help="regex for ncu's kernel name (no template brackets)")
    ap.add_argument("--mangled", default=None, help="substring of the mangled name in the cubin (default: the regex)")
    ap.add_argument("--skip", type=int, default=0)
    ap.add_argument("--per", type=float, default=1.0)
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--by-opcode", action="store_true")
    a = ap.parse_args()
    counts = ncu_counts(a.rep, a.kernel, a.skip)
    lines = disasm_lines(a.cubin, a.mangled or a.kernel)
    print(f"# ncu instructions: {len(counts)}, nvdisasm instructions: {len(lines)}")
    n = min(len(counts), len(lines))
    per_line, per_op, samples, total = collections.Counter(), collections.Counter(), collections.Counter(), 0
    for (src, inst, smp), (loc, _txt) in zip(counts[:n], lines[:n]):
        per_line[loc] += inst
        samples[loc] += smp
        tok = src.split()
        per_op[tok[1] if tok and tok[0].startswith("@") and len(tok) > 1 else (tok[0] if tok else "?")] += inst
        total += inst
    print(f"# total warp-instructions {total}  (/{a.per:g} = {total / a.per:.1f})")
    if a.by_opcode:
        for op, c in per_op.most_common(a.top):
            print(f"{c / a.per:10.2f}  {100 * c / total:5.1f}%  {op}")
        return
    for loc, c in sorted(per_line.items(), key=lambda x: (x[0][0], x[0][1])):
        if c / total >= 0.002:
            print(f"{loc[0]}:{loc[1]:<5d} {c / a.per:10.2f}  {100 * c / total:5.1f}%   samples {samples[loc]}")


if __name__ == "__main__":
    main()
