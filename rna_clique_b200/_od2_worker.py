"""Helper process of filtering_step: turns downloaded table columns into the labelled DataFrame and writes
od2/<S1>--<S2>.<ext> (find_all_pairs.py:57-88 does this per pair in the reference's joblib workers).

Why a process: building the frame and pickling it is pandas / pickle code that holds the interpreter lock — seven
threads write 1.5 times the tables of one — and at BASELINE configs[2] the 435 tables were what rna_clique() waited
for.  A handful of these processes, started once and kept with the shared engine, take that work off the lock of the
process that drives the GPU.  They never touch CUDA: the only native code they call is host code (rcq_table_stats).

Protocol (one request at a time per process): a JSON line on stdin
    {"block": file holding int32[15][n] + uint8 minus[n] + uint8 reverse[n], "n": n, "db": [4 ints],
     "t1": path, "t2": path, "out": path}
answered by one line on stdout: "ok" or "err <message>".  The block file is removed by the worker."""
from __future__ import annotations

import json
import os
import sys


def main() -> None:
    out = sys.stdout
    sys.stdout = sys.stderr                      # nothing but the protocol goes to the real stdout
    import numpy as np
    from pathlib import Path
    from . import _native
    from .find_all_pairs import _label_and_save
    from .find_homologs import table_to_frame
    names = list(_native.TABLE_COLS)
    out.write("ready\n")
    out.flush()
    for line in sys.stdin:
        line = line.strip()
        if not line:
            continue
        try:
            req = json.loads(line)
            n = int(req["n"])
            raw = np.fromfile(req["block"], dtype=np.uint8)
            os.remove(req["block"])
            k = len(names)
            assert raw.size == 4 * k * n + 2 * n, "block file has the wrong size"
            block = raw[:4 * k * n].view(np.int32).reshape(k, n)
            cols = {c: block[i] for i, c in enumerate(names)}
            cols["minus"] = raw[4 * k * n:4 * k * n + n]
            cols["reverse"] = raw[4 * k * n + n:]
            _label_and_save(table_to_frame(cols, *(int(x) for x in req["db"])), Path(req["t1"]), Path(req["t2"]), Path(req["out"]))
            out.write("ok\n")
        except BaseException as e:  # noqa: BLE001 - reported to the parent, which raises it
            out.write("err " + repr(e).replace("\n", " ") + "\n")
        out.flush()


if __name__ == "__main__":
    main()
