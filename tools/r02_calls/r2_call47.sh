#!/bin/bash
# Round 2, call 47 (1 GPU, the round's last GPU seconds): smoke() with its new EAP leg
mkdir -p gpurun_out
O=gpurun_out/c47
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > "$O.smoke.log" 2>&1
echo "rc=$? smoke" | tee "$O.rc"
tail -n 4 "$O.smoke.log"
