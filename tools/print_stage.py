#!/usr/bin/env python
"""print value / ms per step / sample-stage times of a bench.py JSON line (helper of the A/B scripts)"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["config"].get("sample_type"), f"{d['value'] / 1e9:.3f} G/s", f"{d['ms_per_step']:.4f} ms",
      {k: v["ms"] for k, v in d["stages"].items() if k.startswith("sample")})
