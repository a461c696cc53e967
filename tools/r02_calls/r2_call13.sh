#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/c13
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run transport 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x
run tests 1200 python -m pytest tests -x -q -m gpu
tail -n 3 "$O.transport.log"; tail -n 3 "$O.tests.log"
cat "$O.rc"
