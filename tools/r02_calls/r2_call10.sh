#!/bin/bash
# Round 2, call 10 (2 GPUs): the ghost mailbox (flag-in-data E-W exchange) with real peers -- parity, then timing
mkdir -p gpurun_out
O=gpurun_out/c10
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
B="--no-cpu-baseline --no-e2e --no-other-configs --steps 10"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
run parity_tp1 300 $TR 29541 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 3
run parity_tp1_full 300 $TR 29542 tests/run_multigpu_parity.py --grid tp1 --scenario full --calls 2 --ndte 7
run parity_gx1 300 $TR 29543 tests/run_multigpu_parity.py --grid gx1 --scenario full --calls 2
run s8_n2 400 $TR 29544 bench.py --gpus 2 $B --grid tp01s8
run tp01_n2 600 $TR 29545 bench.py --gpus 2 $B
run tenday_tp1_n2 600 $TR 29546 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 60 --evolve 1 --out gpurun_out/tenday60_tp1_n2_ll.json
grep -h "PARITY\|TENDAY" "$O".*.log | cut -c1-300
grep -h '^{' "$O".s8*.log "$O".tp01*.log | cut -c1-260
cat "$O.rc"
