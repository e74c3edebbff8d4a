#!/usr/bin/env python3
"""Stall samples of the last kernel in an .ncu-rep, bucketed into contiguous SASS regions of `step` instructions."""
import csv
import subprocess
import sys

KEYS = ["stall_long_sb", "stall_short_sb", "stall_wait", "stall_not_selected", "stall_selected", "stall_barrier", "stall_mio",
        "stall_math", "stall_dispatch", "stall_branch_resolving", "stall_no_inst", "stall_lg", "stall_membar", "stall_sleeping"]


def main(path, step=200):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, body = None, []
    for r in rows:
        if r and r[0] == "Address":
            hdr, body = r, []
        elif hdr and len(r) == len(hdr):
            body.append(r)
    ix = {h: i for i, h in enumerate(hdr)}
    I = lambda r, k: int(float(r[ix[k]] or 0)) if k in ix else 0
    tot = sum(I(r, "# Samples") for r in body)
    print("instr %d samples %d" % (len(body), tot))
    print("%-11s %6s %6s %8s  " % ("region", "smpl", "%", "exec/wt") + " ".join("%6s" % k[6:12] for k in KEYS))
    for a in range(0, len(body), step):
        seg = body[a:a + step]
        s = sum(I(r, "# Samples") for r in seg)
        ex = sum(I(r, "Instructions Executed") for r in seg)
        print("%4d-%-6d %6d %5.1f%% %8.0f  " % (a, a + len(seg), s, 100.0 * s / tot, ex / 32768.0) +
              " ".join("%6d" % sum(I(r, k) for r in seg) for k in KEYS))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 200)
