#!/bin/bash
# round 2, call 27: flux / update kernels with chunked loads (2, 4, 8, 11 tracers in flight)
mkdir -p gpurun_out
O=gpurun_out/c27
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k remap > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 2 > $O.fc4.log 2>&1; echo "rc=$? fc4" | tee -a $O.rc
tail -n 2 $O.fc4.log | head -n 1
for v in fc2 fc8 fc11; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 900 python tools/profile_remap.py tp01 2 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 2 $O.$v.log | head -n 1
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"edge_flux|update_fields" --csv --log-file $O.launches.csv \
    python tools/profile_remap.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
grep -v "^==" $O.launches.csv | cut -d, -f5,15- | tail -n 2
