#!/usr/bin/env python3
"""Per-instruction stall summary from `ncu -i X.ncu-rep --page source --csv` (SASS view)."""
import csv
import subprocess
import sys


def main(path, key="stall_long_sb", n=30):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    # the page may hold several kernels, each with its own header line
    hdr, body = None, []
    for r in rows:
        if r and r[0] == "Address":
            hdr, body = r, []  # keep the last kernel
        elif hdr and len(r) == len(hdr):
            body.append(r)
    ix = {h: i for i, h in enumerate(hdr)}
    I = lambda r, k: int(float(r[ix[k]] or 0))
    print("instructions", len(body), "samples", sum(I(r, "# Samples") for r in body),
          "warp-instr executed", sum(I(r, "Instructions Executed") for r in body))
    for k in ("stall_long_sb", "stall_short_sb", "stall_wait", "stall_not_selected", "stall_selected", "stall_barrier", "stall_mio",
              "stall_math", "stall_dispatch", "stall_lg", "stall_branch_resolving", "stall_no_inst", "stall_misc"):
        print("  %-24s %d" % (k, sum(I(r, k) for r in body)))
    print("--- top", key)
    for r in sorted(body, key=lambda r: -I(r, key))[:n]:
        print("%6d %8d  %s" % (I(r, key), I(r, "Instructions Executed"), r[ix["Source"]].strip()[:110]))


if __name__ == "__main__":
    main(sys.argv[1], *(sys.argv[2:3] or ["stall_long_sb"]), *( [int(sys.argv[3])] if len(sys.argv) > 3 else []))
