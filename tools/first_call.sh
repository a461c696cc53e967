#!/bin/bash
# The first 1-GPU box call of a round: everything that was written without hardware gets its first run,
# the default path is re-measured, and the prepared kernel variants are A/B-ed.  Results under gpurun_out/.
#
#   here, first:  python -m cice5_b200.build --variant mergeuv -DEVP_TMA_MERGE_UV=1
#                 python -m cice5_b200.build --variant staticfirst -DEVP_TMA_STATIC_FIRST=1
#                 python -m cice5_b200.build --variant r2all -DEVP_TMA_MERGE_UV=1 -DEVP_TMA_STATIC_FIRST=1
#                 python -m cice5_b200.build --variant guided -DEVP_TMA_GUIDED=1
#                 python -m cice5_b200.build --variant r2guided -DEVP_TMA_MERGE_UV=1 -DEVP_TMA_STATIC_FIRST=1 -DEVP_TMA_GUIDED=1
#   then:         /usr/local/graft/bin/gpurun --timeout 2700 -- 'bash tools/first_call.sh'   (about 40 GPU-minutes; the steps are ordered by importance)
#   afterwards:   python tools/first_call_report.py
#
# Every step runs under its own timeout so that one hang cannot take the box (a strike) or the other steps.
mkdir -p gpurun_out
O=gpurun_out/first
run() { # name seconds command...
  local n=$1 t=$2; shift 2
  echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"
}
: > "$O.rc"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$O.gpu.log" 2>&1

# 1. the validated suite alone (what the round-end tier must keep green), stopping at the first failure
run tests_validated 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_revised_evp.py \
    --deselect tests/test_gpu_strength.py --deselect tests/test_gpu_zw_reference_text.py \
    --deselect tests/test_gpu_zx_wet_pole.py --deselect tests/test_gpu_zy_fullsize.py \
    --deselect tests/test_gpu_zz_c_host.py --deselect tests/test_gpu_zv_pinned_host.py
# 2. the rows whose first hardware run is pending, one file per process (a sticky CUDA error in one
#    cannot poison the next), xfail marks ignored so that a failure shows its traceback
for f in zw_reference_text zx_wet_pole revised_evp strength zz_c_host zv_pinned_host zy_fullsize; do
  run pending_$f 600 python -m pytest tests/test_gpu_$f.py -m gpu -q --runxfail -rA
done
# 3. the default bench line (also the reference arm, for the ratio)
run bench 400 python bench.py
run bench_reference 300 python bench.py --impl reference --steps 2 --warmup 1
# 4. prepared variants: parity on the tripole cases first, then the kernel-only bench
for v in mergeuv staticfirst r2all; do
  lib=cice5_b200/lib/exp_$v.so
  [ -f "$lib" ] || continue
  EVP_B200_LIB=$lib run parity_$v 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tp1 or tripole or gx3"
  EVP_B200_LIB=$lib run bench_$v 300 python bench.py --no-cpu-baseline --no-e2e
done
# 4b. the in-kernel tripole fold (EVP_B200_FOLD=kernel, one launch per subcycle): parity, then the bench
EVP_B200_FOLD=kernel run parity_foldkernel 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tp1 or tripole"
EVP_B200_FOLD=kernel run bench_foldkernel 300 python bench.py --no-cpu-baseline --no-e2e
[ -f cice5_b200/lib/exp_r2all.so ] && EVP_B200_FOLD=kernel EVP_B200_LIB=cice5_b200/lib/exp_r2all.so \
  run bench_r2all_foldkernel 300 python bench.py --no-cpu-baseline --no-e2e
# 4d. guided item sizes (-DEVP_TMA_GUIDED=1 libraries with EVP_B200_GUIDED="percent,rows": short items at the
#     end of the work list): parity, then the bench, alone and on top of the other two variants
for v in guided r2guided; do
  lib=cice5_b200/lib/exp_$v.so
  [ -f "$lib" ] || continue
  EVP_B200_GUIDED=30,3 EVP_B200_LIB=$lib run parity_$v 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tp1 or tripole"
  EVP_B200_GUIDED=30,3 EVP_B200_LIB=$lib run bench_$v 300 python bench.py --no-cpu-baseline --no-e2e
done
# 4c. early download of the outputs that are final before the loop (EVP_B200_EARLY_D2H=1): parity, then e2e
EVP_B200_EARLY_D2H=1 run parity_earlyd2h 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "gx3 or tp1"
EVP_B200_EARLY_D2H=1 run bench_earlyd2h 300 python bench.py --no-cpu-baseline
# 5. launch list of the default bench (the kernel's share of the step), never a bench value
run ncu_launches 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file "$O.launches.csv" python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
grep -h '^{' "$O".bench*.log | cut -c1-400
cat "$O.rc"
