#!/bin/bash
# Round 2, call 2 (1 GPU): first hardware run of the persistent loop kernel (evp_loop_tma_kernel).
mkdir -p gpurun_out
O=gpurun_out/c2
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run smoke 300 python __graft_entry__.py --smoke
run tests 900 python -m pytest tests -m gpu -q -x
run bench 300 python bench.py --no-cpu-baseline
EVP_B200_LOOP=0 run bench_noloop 300 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs
for g in tp01s8 gx3 gx1 tp1 tp025; do
  run bench_$g 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs --grid $g --steps 10
done
for rpc in 4 8 12 16 24 32; do
  EVP_B200_RPC=$rpc run bench_s8_rpc$rpc 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs --grid tp01s8 --steps 10
done
for rpc in 8 16 48 64; do
  EVP_B200_RPC=$rpc run bench_rpc$rpc 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs
done
for rpc in 2 3 8; do
  EVP_B200_RPC=$rpc run bench_gx1_rpc$rpc 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs --grid gx1 --steps 10
  EVP_B200_RPC=$rpc run bench_tp1_rpc$rpc 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs --grid tp1 --steps 10
done
run bench_caps 200 python bench.py --no-cpu-baseline --no-e2e --no-parity --no-other-configs --scenario caps
# memcheck on the small grids (one sanitizer tool per call)
for k in tma sweep; do
  run san_memcheck_gx3_$k 300 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_case.py gx3 $k 3
  run san_memcheck_tp_$k 300 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_case.py tp96x40 $k 3
done
run san_memcheck_tp_tma_r2 300 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_case.py tp96x40 tma 3 2
grep -h '^{' "$O".bench*.log | cut -c1-200
cat "$O.rc"
