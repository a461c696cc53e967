#!/bin/bash
# Round 2, call 9 (1 GPU): the allocation-free entry point -- alone in a fresh process, then after the whole suite
mkdir -p gpurun_out
O=gpurun_out/c9
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run noalloc_alone 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x
run tests 1200 python -m pytest tests -m gpu -q -x
tail -3 "$O.noalloc_alone.log" "$O.tests.log"
cat "$O.rc"
