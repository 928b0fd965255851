#!/usr/bin/env python
"""Prints the headline fields of one or more bench.py JSON lines."""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f)
    for k in ("value", "ms_per_step", "extension_steps_per_step", "gpu_launches", "seconds_to_matrix"):
        print("  ", k, d.get(k))
    if d.get("e2e"):
        print("   e2e", {k: v for k, v in d["e2e"].items() if k != "what"})
    for name in ("roofline", "roofline_ext"):
        r = d.get(name)
        if r:
            print("  ", name, {k: r[k] for k in ("kernel", "achieved", "peak", "frac", "launch_ms") if k in r},
                  {k: r[k] for k in ("extension_steps_per_launch", "gcups_alone") if k in r})
    if d.get("cpu_baseline"):
        print("   cpu", {k: v for k, v in d["cpu_baseline"].items() if k != "sample"})
    print("   host_ms", d.get("host_ms_last_step"), "clocks", d.get("clocks"))
