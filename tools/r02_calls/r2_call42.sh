#!/bin/bash
# Round 2, call 42 (1 GPU): EAP at 0.1 degree -- bitwise against the oracle at tp01 (ndte = 11), device time per subcycle,
# launch list and one `ncu --set full` capture of eap_subcycle_kernel
mkdir -p gpurun_out
O=gpurun_out/c42
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run eap_tp01 150 python tools/profile_eap.py tp01 11 3 1
run launches 90 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file "$O.launches.csv" python tools/profile_eap.py tp01 11 1 0
run ncu_full 120 ncu --set full --clock-control none --import-source on -k regex:eap_subcycle_kernel -s 3 -c 1 -o "$O.eap_full" -f python tools/profile_eap.py tp01 11 1 0
tail -n 8 "$O.eap_tp01.log"
cat "$O.rc"
