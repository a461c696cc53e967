#!/bin/bash
# round 2, call 14: HEAD after the source clean-up + field reconstruction: whole GPU tier, smoke, default bench line
mkdir -p gpurun_out
O=gpurun_out/c14
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run transport 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x
run tests 1200 python -m pytest tests -x -q -m gpu
run smoke 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
run bench 900 python bench.py --steps 5 --warmup 3
tail -n 3 "$O.transport.log"; tail -n 3 "$O.tests.log"; tail -n 2 "$O.smoke.log"
python - <<'PY'
import json
for l in open("gpurun_out/c14.bench.log"):
    if l.startswith("{"):
        r = json.loads(l)
        print("value", r["value"], "frac", r["roofline"]["frac"], "e2e", r["e2e"]["value"] if r.get("e2e") else None)
        print("transport", json.dumps(r.get("transport")))
PY
cat "$O.rc"
