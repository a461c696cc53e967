#!/bin/bash
# Round 2, call 45 (2 GPUs): EAP with one process per GPU (CUDA IPC arenas, the cross-rank flag barrier between the stages
# of a subcycle, real concurrency), bitwise against the one-block oracle; then the device EAP tests on the final library
mkdir -p gpurun_out
O=gpurun_out/c45
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run eap_2proc 70 python tests/run_multigpu_eap.py --nranks 2 --grid tp1 --scenario caps --ndte 21 --calls 2
EVP_TEST_CHILD=1 run eap 60 python -m pytest tests/test_zy_gpu_eap.py -m gpu -q -x -s
tail -n 3 "$O.eap_2proc.log" "$O.eap.log"
cat "$O.rc"
