#!/bin/bash
# Round 2, call 41 (1 GPU, the last GPU minutes of the round): EAP on hardware -- the three device tests of
# tests/test_zy_gpu_eap.py in one process (prints the per-subcycle device time), then the allocation-free entry point
mkdir -p gpurun_out
O=gpurun_out/c41
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
EVP_TEST_CHILD=1 run eap 200 python -m pytest tests/test_zy_gpu_eap.py -m gpu -q -x -s
run noalloc 100 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x
tail -6 "$O.eap.log" "$O.noalloc.log"
cat "$O.rc"
