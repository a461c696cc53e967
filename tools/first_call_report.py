#!/usr/bin/env python
"""Summary of a tools/first_call.sh run: exit codes, the last line of every pytest log, the bench lines."""
import glob
import json
import os
import sys

d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
rc = os.path.join(d, "first.rc")
if os.path.exists(rc):
    print(open(rc).read().strip())
for log in sorted(glob.glob(os.path.join(d, "first.*.log"))):
    lines = [l.rstrip() for l in open(log, errors="replace") if l.strip()]
    name = os.path.basename(log)[len("first."):-len(".log")]
    if name.startswith(("tests_", "pending_", "parity_")):
        print(f"{name:28s} {lines[-1] if lines else '(empty)'}")
        for l in lines:
            if l.startswith(("FAILED", "ERROR")):
                print(" " * 29 + l[:200])
    elif name.startswith("bench"):
        for l in lines:
            if l.startswith("{"):
                j = json.loads(l)
                roof = j.get("roofline") or {}
                e2e = j.get("e2e") or {}
                print(f"{name:28s} value {j.get('value'):.4g} {j.get('unit', '')}  ms/step {j.get('ms_per_step')}"
                      f"  frac {roof.get('frac')}  e2e {e2e.get('value')}  clocks {j.get('clocks')}")
