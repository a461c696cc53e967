#!/bin/bash
# Round 2, call 46 (1 GPU, what is left of the budget): bench.py with the new `eap` block on the 1-degree configuration
mkdir -p gpurun_out
O=gpurun_out/c46
timeout 95 python bench.py --grid tp1 --no-cpu-baseline --no-e2e --no-other-configs --steps 3 --warmup 3 > "$O.bench_tp1.json" 2> "$O.bench_tp1.err"
echo "rc=$? bench_tp1" | tee "$O.rc"
tail -c 1500 "$O.bench_tp1.json"; tail -n 3 "$O.bench_tp1.err"
