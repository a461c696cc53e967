#!/bin/bash
# round 2, call 39: horizontal_remap on tp01 with the "caps" ice cover (a third of the cells hold ice)
mkdir -p gpurun_out
O=gpurun_out/c39
timeout 900 python tools/profile_remap.py tp01 2 1 caps > $O.caps.log 2>&1; echo "rc=$? caps" | tee $O.rc
tail -n 3 $O.caps.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"edge_flux|update_fields|reconstruct" --csv --log-file $O.launches.csv \
    python tools/profile_remap.py tp01 1 1 caps > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
grep -v "^==" $O.launches.csv | cut -d, -f5,13- | tail -n 9
