#!/usr/bin/env python
"""Table of the per-rank device timeline lines that bench.py prints with EVP_B200_TRACE=1."""
import re
import sys

for path in sys.argv[1:]:
    txt = open(path).read()
    rows = re.findall(r"\[trace rank (\d+)\][^:]*: kernel ([-\d.]+)\s+signal ([-\d.]+)\s+to-fold ([-\d.]+)\s+"
                      r"wait ([-\d.]+)\s+fold\+next-launch ([-\d.]+)\s+period ([-\d.]+)", txt)
    print(f"## {path}\n| rank | kernel | signal | to-fold | wait | fold+launch | period |\n|---:|---:|---:|---:|---:|---:|---:|")
    for r in sorted(rows, key=lambda x: int(x[0])):
        print("| " + " | ".join(r) + " |")
