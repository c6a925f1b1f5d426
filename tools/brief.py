#!/usr/bin/env python
"""Condense bench.py JSON lines (stdin) into one short line each; argv[1:] is a free-form tag."""
import json
import sys

tag = " ".join(sys.argv[1:])
for line in sys.stdin:
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    if "rays" not in d:
        print(tag, d.get("impl"), d.get("value"))
        continue
    r = d["rays"]
    rb = d.get("roofline_brute") or {}
    print(tag, d["config"].get("pipeline"), d["config"].get("accel"), "n_gpus", d["n_gpus"], "ms %.2f" % d["ms_per_step"],
          "Mrays/s %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], "exact", r.get("exact_tests"), "nodes", r.get("node_visits"),
          "frac %.3f" % (d["roofline"]["frac"] or 0), "brute_frac %.3f ms %.1f" % (rb.get("frac", 0), rb.get("ms", 0)))
