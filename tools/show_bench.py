#!/usr/bin/env python
"""Print the essentials of bench.py JSON lines (files given on the command line)."""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.3f value %.3g e2e %.3g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]))
        print("  stages", {k: round(v, 3) for k, v in d["stage_ms_per_step"].items()}, "colors %.1f" % d["config"].get("avg_colors", 0),
              "checksum", d["e2e"].get("checksum"))
    except Exception as e:
        print(f, "ERR", e)
