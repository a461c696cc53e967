#!/bin/bash
# Round 2, call 5 (2 GPUs, one process per GPU): the loop kernel with real peers -- parity, the 10-day run, bench
mkdir -p gpurun_out
O=gpurun_out/c5
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
run parity_tp1 300 $TR 29511 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 2
run parity_tp96 300 $TR 29512 tests/run_multigpu_parity.py --grid tp1 --scenario full --calls 3 --ndte 7
run parity_gx1 300 $TR 29513 tests/run_multigpu_parity.py --grid gx1 --scenario full --calls 2
run bench2 600 $TR 29514 bench.py --gpus 2
run tenday_tp1_n2 900 $TR 29515 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 240 --evolve 1 --out gpurun_out/tenday_tp1_n2.json
run tenday_tp1_n1 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29516 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 240 --evolve 1 --out gpurun_out/tenday_tp1_n1.json
run tenday_tp025_n2 900 $TR 29517 tests/run_multigpu_parity.py --grid tp025 --scenario caps --calls 12 --evolve 1 --out gpurun_out/tenday_tp025_n2.json
EVP_B200_LOOP=0 run bench2_noloop 600 $TR 29518 bench.py --gpus 2 --no-e2e --no-parity --no-other-configs
grep -h "PARITY\|TENDAY" "$O".*.log | cut -c1-400
grep -h '^{' "$O".bench*.log | cut -c1-300
cat "$O.rc"
