#!/bin/bash
# statevector gate engine: timing at the C5 shape, then one ncu --set full of sv_group_kernel
mkdir -p gpurun_out
timeout 900 python scripts/time_uncut.py 29 1 > gpurun_out/sv_time_n29.log 2>&1; echo "n29 rc=$?"; cat gpurun_out/sv_time_n29.log | tail -8
timeout 600 python scripts/time_uncut.py 25 2 > gpurun_out/sv_time_n25.log 2>&1; echo "n25 rc=$?"; cat gpurun_out/sv_time_n25.log | tail -8
cat > /tmp/one_case.py <<'P'
import sys
sys.argv = ["x"]
sys.path.insert(0, "scripts")
import time_uncut
time_uncut.case(27, 1, 12, 4)
P
timeout 600 python /tmp/one_case.py > gpurun_out/sv_one_case.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:sv_group -s 4 -c 4 \
   -o gpurun_out/r2_sv_group_n28 python /tmp/one_case.py > gpurun_out/sv_ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/sv_ncu.log
ls -la gpurun_out/*.ncu-rep
