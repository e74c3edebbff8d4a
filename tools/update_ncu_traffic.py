#!/usr/bin/env python3
"""Rewrite bench.py's NCU_TRAFFIC numbers from profiles/r02_<name>_ncu_full_summary.txt (dram__bytes_read/write per launch)."""
import os
import re
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def traffic(name):
    txt = open(os.path.join(REPO, "profiles", "r02_%s_ncu_full_summary.txt" % name)).read()
    out = []
    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        m = re.search(re.escape(key) + r"\s+([0-9.]+)\s+(\w+)", txt)
        out.append(float(m.group(1)) * UNIT[m.group(2)])
    return out


def main():
    path = os.path.join(REPO, "bench.py")
    src = open(path).read()

    def repl(m):
        r, w = traffic(m.group(1))
        return '_t("%s", %.6fe6, %.6fe6)' % (m.group(1), r / 1e6, w / 1e6)

    new = re.sub(r'_t\("(\w+)", [0-9.e]+, [0-9.e]+\)', repl, src)
    open(path, "w").write(new)
    print("updated" if new != src else "unchanged")


if __name__ == "__main__":
    sys.exit(main())
