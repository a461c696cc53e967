#!/bin/bash
# Round 2, call 7 (8 GPUs): strong scaling of tp01 at N = 8 and N = 4, parity leg at both
mkdir -p gpurun_out
O=gpurun_out/c7
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run bench8 600 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8
run bench4 600 $TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --no-other-configs
run parity8_tp1 300 $TR --nproc-per-node 8 --master-port 29523 tests/run_multigpu_parity.py --grid tp1 --scenario caps --calls 2
EVP_B200_LOOP=0 run bench8_noloop 400 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --no-e2e --no-parity --no-other-configs
grep -h "PARITY" "$O".*.log | cut -c1-300
grep -h '^{' "$O".bench*.log | cut -c1-300
cat "$O.rc"
