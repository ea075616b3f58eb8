#!/usr/bin/env python3
"""One-line digest of bench.py JSON lines: tools/jl.py file.json [...]"""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read())
    except Exception as e:
        print(f, "unreadable:", e); continue
    print(f, round(d["value"]), {k: round(v, 2) for k, v in d.get("stages_ms", {}).items()}, "msm_ms", round(d["roofline"]["launch_ms"], 2), "frac", round(d["roofline"]["frac"], 3),
          "e2e", round(d["e2e"]["value"]), round(d["e2e"].get("ms_per_step", 0), 1))
