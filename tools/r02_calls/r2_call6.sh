#!/bin/bash
# Round 2, call 6 (1 GPU): mbarrier wait with a suspend-time hint vs spinning; full bench line with the new e2e legs
mkdir -p gpurun_out
O=gpurun_out/c6
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
B="--no-cpu-baseline --no-e2e --no-parity --no-other-configs"
run tests 1200 python -m pytest tests -m gpu -q -x
run bench 300 python bench.py $B
EVP_B200_LIB=cice5_b200/lib/exp_nohint.so run bench_nohint 300 python bench.py $B
run bench_s8 200 python bench.py $B --grid tp01s8 --steps 10
EVP_B200_LIB=cice5_b200/lib/exp_nohint.so run bench_s8_nohint 200 python bench.py $B --grid tp01s8 --steps 10
run bench_caps 200 python bench.py $B --scenario caps
EVP_B200_LOOP=0 run bench_noloop 300 python bench.py $B
EVP_B200_LOOP=0 EVP_B200_LIB=cice5_b200/lib/exp_nohint.so run bench_noloop_nohint 300 python bench.py $B
run bench_full 900 python bench.py
python bench.py $B --steps 1 --warmup 1 > "$O.plain.log" 2>&1 && \
run ncu_full 900 ncu --set full --import-source on --clock-control none -k regex:evp_loop_tma_kernel -s 1 -c 1 \
    -o "$O.loop_full" -f python bench.py $B --steps 1 --warmup 1
grep -h '^{' "$O".bench*.log | cut -c1-200
cat "$O.rc"
