#!/bin/bash
# Round 2, call 44 (1 GPU, the round's last GPU minute): what the round-end tier runs first -- smoke() -- and the two newest
# device test files through the ordinary pytest path (child processes for the `isolated` EAP tests)
mkdir -p gpurun_out
O=gpurun_out/c44
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run smoke 100 python -c "import __graft_entry__ as g; g.smoke()"
run newest 100 python -m pytest tests/test_zy_gpu_eap.py tests/test_zz_no_alloc.py tests/test_gpu_revised_evp.py -m gpu -q -x
tail -n 4 "$O.smoke.log" "$O.newest.log"
cat "$O.rc"
