#!/usr/bin/env python3
"""Rewrite the size-sweep table of DESIGN.md from a tools/sweep.py JSONL file (cases h and c)."""
import json
import re
import sys

src = sys.argv[1]
rows = {}
for l in open(src):
    d = json.loads(l)
    rows.setdefault(d["envs"], {})[d["case"]] = d
out = ["| envs | humanoid3d walk µs | env-steps/s | HBM frac | G1 combined µs | env-steps/s | HBM frac |", "|---|---|---|---|---|---|---|"]
for n in sorted(rows):
    h, c = rows[n]["h"], rows[n]["c"]
    out.append("| %d | %.1f | %.2e | %.3f | %.1f | %.2e | %.3f |" % (n, h["ms"] * 1e3, h["env_steps_per_s"], h["frac"], c["ms"] * 1e3,
                                                                    c["env_steps_per_s"], c["frac"]))
p = "DESIGN.md"
s = open(p).read()
pat = re.compile(r"\| envs \| humanoid3d walk µs .*?\n\n", re.S)
assert pat.search(s)
s = pat.sub("\n".join(out) + "\n\n", s, count=1)
s = re.sub(r"`profiles/r01_sweep_1gpu[_a-z0-9]*\.jsonl`\), statistics on", "`%s`), statistics on" % src, s)
open(p, "w").write(s)
