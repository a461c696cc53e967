#!/bin/bash
# round 2, call 24: HEAD with the final reconstruction kernel: GPU tier, bench line, ncu of reconstruct_kernel
mkdir -p gpurun_out
O=gpurun_out/c24
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run tests 1200 python -m pytest tests -x -q -m gpu
run bench 900 python bench.py --steps 5 --warmup 3
run plain 600 python tools/profile_reconstruct.py tp01 3
run ncu 900 ncu --set full --clock-control none --import-source on -k regex:reconstruct_kernel -c 1 -o $O.reconstruct -f python tools/profile_reconstruct.py tp01 1
tail -n 3 "$O.tests.log"; cat $O.plain.log
python - <<'PY'
import json
for l in open("gpurun_out/c24.bench.log"):
    if l.startswith("{"):
        r = json.loads(l)
        print("value", r["value"], "frac", r["roofline"]["frac"], "e2e", r["e2e"]["value"] if r.get("e2e") else None)
        print("transport", json.dumps(r.get("transport")))
PY
cat "$O.rc"
