#!/bin/bash
# Round 2, call 4 (1 GPU): loop kernel v2 (ndte-1 subcycles, one-thread publish, gpu-scope counters, prefetched queue
# position, corner-wise stress_cell) against a launch per subcycle; e2e with the resident protocol
mkdir -p gpurun_out
O=gpurun_out/c4
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
B="--no-cpu-baseline --no-e2e --no-parity --no-other-configs"
run tests 1200 python -m pytest tests -m gpu -q -x
run bench 300 python bench.py $B
EVP_B200_LOOP=0 run bench_noloop 300 python bench.py $B
for v in lx3 lx7; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so run bench_$v 200 python bench.py $B
  EVP_B200_RPC=8 EVP_B200_LIB=cice5_b200/lib/exp_$v.so run bench_s8_$v 200 python bench.py $B --grid tp01s8 --steps 10
done
for rpc in 4 8 12 16 20; do
  EVP_B200_RPC=$rpc run bench_s8_rpc$rpc 200 python bench.py $B --grid tp01s8 --steps 10
done
EVP_B200_LOOP=0 run bench_s8_noloop 200 python bench.py $B --grid tp01s8 --steps 10
for rpc in 16 24 48; do
  EVP_B200_RPC=$rpc run bench_rpc$rpc 200 python bench.py $B
done
for g in gx3 gx1 tp1 tp025; do
  run bench_$g 200 python bench.py $B --grid $g --steps 10
  EVP_B200_LOOP=0 run bench_${g}_noloop 200 python bench.py $B --grid $g --steps 10
done
EVP_B200_RPC=2 run bench_tp1_rpc2 200 python bench.py $B --grid tp1 --steps 10
EVP_B200_RPC=2 run bench_gx1_rpc2 200 python bench.py $B --grid gx1 --steps 10
run bench_full 600 python bench.py
python bench.py $B --steps 1 --warmup 1 > "$O.plain.log" 2>&1 && \
run ncu_full 900 ncu --set full --import-source on --clock-control none -k regex:evp_loop_tma_kernel -s 1 -c 1 \
    -o "$O.loop_full" -f python bench.py $B --steps 1 --warmup 1
grep -h '^{' "$O".bench*.log | cut -c1-200
cat "$O.rc"
