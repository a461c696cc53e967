#!/bin/bash
# Round 2, call 12 (1 GPU): what the driver runs at round end -- device suite, smoke, the bench as the driver calls it,
# the reference arm -- plus the ncu launch list of a short bench run
mkdir -p gpurun_out
O=gpurun_out/c12
run() { local n=$1 t=$2; shift 2; echo "== $n"; timeout "$t" "$@" > "$O.$n.log" 2>&1; echo "rc=$? $n" | tee -a "$O.rc"; }
: > "$O.rc"
run tests 1200 python -m pytest tests -x -q -m gpu
run smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
run bench_driver 900 python bench.py --gpus 1 --steps 20 --warmup 5
run bench_reference 400 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1
B="--no-cpu-baseline --no-e2e --no-parity --no-other-configs --steps 2 --warmup 1"
python bench.py $B > "$O.plain.log" 2>&1 && \
run ncu_launches 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file "$O.launches.csv" python bench.py $B
tail -n 2 "$O.tests.log"; tail -n 1 "$O.smoke.log"
grep -h '^{' "$O".bench*.log | cut -c1-300
cat "$O.rc"
