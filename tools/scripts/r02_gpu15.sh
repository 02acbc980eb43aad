#!/bin/bash
# Round 2, GPU call 15: ncu --set full of dwconv7_v3 (warp-decoupled) at 64x56x56x96, source counters.
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
CPN_DW_IMPL=v3 timeout 300 ncu --set full --clock-control none --import-source on -k regex:dwconv7_v3 --launch-skip 2 -c 1 -o /tmp/ncu_r02/dw3b -f python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 3 > gpurun_out/r02_15_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/ncu_r02/dw3b.ncu-rep --page raw --csv > gpurun_out/r02_15_ncu_dw3_raw.csv 2>/dev/null
ncu -i /tmp/ncu_r02/dw3b.ncu-rep --page source --csv > gpurun_out/r02_15_ncu_dw3_source.csv 2>/dev/null
ls -la gpurun_out/r02_15_ncu*
