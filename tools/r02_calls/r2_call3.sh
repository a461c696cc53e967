#!/bin/bash
# Round 2, call 3 (1 GPU): where does the persistent loop kernel lose time?  timing experiments + ncu --set full
mkdir -p gpurun_out
O=gpurun_out/c3
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
B="--no-cpu-baseline --no-e2e --no-parity --no-other-configs"
run tests 1200 python -m pytest tests -m gpu -q -x
run bench 300 python bench.py $B
for v in lx8 lx9 lx11 lx15; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so run bench_$v 200 python bench.py $B
  EVP_B200_RPC=8 EVP_B200_LIB=cice5_b200/lib/exp_$v.so run bench_s8_$v 200 python bench.py $B --grid tp01s8 --steps 10
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so run bench_tp1_$v 200 python bench.py $B --grid tp1 --steps 10
done
EVP_B200_RPC=8 run bench_s8 200 python bench.py $B --grid tp01s8 --steps 10
run bench_e2e 300 python bench.py --no-cpu-baseline --no-parity --no-other-configs
python bench.py $B --steps 1 --warmup 1 > "$O.plain.log" 2>&1 && \
run ncu_full 900 ncu --set full --import-source on --clock-control none -k regex:evp_loop_tma_kernel -s 1 -c 1 \
    -o "$O.loop_full" -f python bench.py $B --steps 1 --warmup 1
grep -h '^{' "$O".bench*.log | cut -c1-200
cat "$O.rc"
