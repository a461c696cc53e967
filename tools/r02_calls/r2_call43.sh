#!/bin/bash
# Round 2, call 43 (1 GPU): the two-stage EAP subcycle (eap_yield_kernel: one thread per (cell, corner); then the row sweep)
# -- device tests, 0.1 degree bitwise + timing, launch list, one `ncu --set full` capture of eap_yield_kernel
mkdir -p gpurun_out
O=gpurun_out/c43
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
EVP_TEST_CHILD=1 run eap 120 python -m pytest tests/test_zy_gpu_eap.py -m gpu -q -x -s
run eap_tp01 120 python tools/profile_eap.py tp01 11 3 1
run launches 80 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file "$O.launches.csv" python tools/profile_eap.py tp01 11 1 0
run ncu_full 100 ncu --set full --clock-control none --import-source on -k regex:eap_yield_kernel -s 3 -c 1 -o "$O.eap_yield_full" -f python tools/profile_eap.py tp01 11 1 0
tail -n 5 "$O.eap.log" "$O.eap_tp01.log"
cat "$O.rc"
