#!/bin/bash
# Round 2, call 8 (2 GPUs): what does an edge-strip item cost?  tp01s8 cut in two: every strip of both ranks is an edge strip
mkdir -p gpurun_out
O=gpurun_out/c8
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
B="--no-cpu-baseline --no-e2e --no-parity --no-other-configs --steps 10"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
run s16_1gpu 200 python bench.py $B --grid tp01s16
run s8_n2 300 $TR 29531 bench.py --gpus 2 $B --grid tp01s8
P=29532
for v in px16 px32 px48 px112; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so run s8_n2_$v 300 $TR $P bench.py --gpus 2 $B --grid tp01s8
  P=$((P+1))
done
EVP_B200_LOOP=0 run s8_n2_noloop 300 $TR 29540 bench.py --gpus 2 $B --grid tp01s8
for rpc in 6 24; do
  EVP_B200_RPC=$rpc run s8_n2_rpc$rpc 300 $TR $P bench.py --gpus 2 $B --grid tp01s8
  P=$((P+1))
done
grep -h '^{' "$O".s*.log | cut -c1-200
cat "$O.rc"
