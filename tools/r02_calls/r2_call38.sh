#!/bin/bash
# round 2, call 38: HEAD with transport_remap: whole GPU tier, smoke, default bench line
mkdir -p gpurun_out
O=gpurun_out/c38
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run tests 1500 python -m pytest tests -x -q -m gpu
run smoke 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
run bench 1200 python bench.py --steps 5 --warmup 3
tail -n 3 "$O.tests.log"; tail -n 2 "$O.smoke.log"
python - <<'PY'
import json
for l in open("gpurun_out/c38.bench.log"):
    if l.startswith("{"):
        r = json.loads(l)
        print("value", r["value"], "frac", r["roofline"]["frac"], "e2e", r["e2e"]["value"] if r.get("e2e") else None)
        print("transport", json.dumps(r.get("transport")))
PY
cat "$O.rc"
