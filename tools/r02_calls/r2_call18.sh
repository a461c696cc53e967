#!/bin/bash
# round 2, call 18: ncu --set full of reconstruct_tma_kernel on tp01 (after the same command ran plainly); batched halos
mkdir -p gpurun_out
O=gpurun_out/c18
timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.plain.log 2>&1; echo "rc=$? plain" | tee $O.rc
timeout 900 ncu --set full --clock-control none --import-source on -k regex:reconstruct_tma_kernel -c 1 -o $O.reconstruct_tma -f \
    python tools/profile_reconstruct.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
timeout 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
cat $O.plain.log; tail -n 3 $O.ncu.log; tail -n 2 $O.transport.log
