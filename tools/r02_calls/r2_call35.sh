#!/bin/bash
# round 2, call 35: flux kernel with the triangles worked out once per edge (warp 0) and shared by the category warps
mkdir -p gpurun_out
O=gpurun_out/c35
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k "remap" > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 2 > $O.fb4.log 2>&1; echo "rc=$? fb4 (product)" | tee -a $O.rc
tail -n 2 $O.fb4.log | head -n 1
for v in fb3 fb5 fc2; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 900 python tools/profile_remap.py tp01 2 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 2 $O.$v.log | head -n 1
done
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"edge_flux|update_fields" --csv --log-file $O.launches.csv \
    python tools/profile_remap.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
grep -v "^==" $O.launches.csv | cut -d, -f5,13- | tail -n 6
