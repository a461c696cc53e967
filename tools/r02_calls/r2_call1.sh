#!/bin/bash
# Round 2, first 1-GPU box call: full device suite without xfail marks, default bench, A/B of the prepared
# kernel variants on tp01 and on the N = 8 per-GPU proxy (tp01s8: 450 x 2700), ncu of the shipped kernel,
# compute-sanitizer on small grids.  Every step under its own timeout.  Results: gpurun_out/c1.*
mkdir -p gpurun_out
O=gpurun_out/c1
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$O.gpu.log" 2>&1
run tests 900 python -m pytest tests -m gpu -q -rA
run bench 300 python bench.py --no-cpu-baseline
run bench_s8 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 10
for v in mergeuv staticfirst r2all stages3 r2s3; do
  lib=cice5_b200/lib/exp_$v.so
  [ -f "$lib" ] || continue
  EVP_B200_LIB=$lib run parity_$v 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tp1 or tripole"
  EVP_B200_LIB=$lib run bench_$v 200 python bench.py --no-cpu-baseline --no-e2e
  EVP_B200_LIB=$lib run bench_s8_$v 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 10
done
EVP_B200_FOLD=kernel run parity_foldkernel 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tp1 or tripole"
EVP_B200_FOLD=kernel run bench_foldkernel 200 python bench.py --no-cpu-baseline --no-e2e
EVP_B200_FOLD=kernel run bench_s8_foldkernel 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 10
EVP_B200_FOLD=kernel EVP_B200_LIB=cice5_b200/lib/exp_r2all.so run bench_s8_r2all_foldkernel 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 10
for rpc in 4 8 16; do
  EVP_B200_RPC=$rpc run bench_s8_rpc$rpc 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 10
done
EVP_B200_TRACE=1 run trace_s8 200 python bench.py --no-cpu-baseline --no-e2e --grid tp01s8 --steps 3
# small configs: per-subcycle cost today
for g in gx3 gx1 tp1 tp025; do
  run bench_$g 200 python bench.py --no-cpu-baseline --no-e2e --grid $g --steps 10
done
# ncu: launch list, then --set full of the shipped kernel (never a bench value)
run ncu_launches 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file "$O.launches.csv" python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
run ncu_full 600 ncu --set full --import-source on --clock-control none -k regex:evp_subcycle_tma_kernel -s 130 -c 2 \
    -o "$O.subcycle_full" -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
grep -h '^{' "$O".bench*.log | cut -c1-300
cat "$O.rc"
