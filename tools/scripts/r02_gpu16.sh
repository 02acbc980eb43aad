#!/bin/bash
# Round 2, GPU call 16: dwconv7_v3 (16 rotating warps): kernel tests with v3 forced, A/B v1 vs v3.
cd /root/repo
mkdir -p gpurun_out
CPN_DW_IMPL=v3 timeout 200 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats" > gpurun_out/r02_16_dw3_tests.log 2>&1
echo "dw3 tests rc=$?"; tail -6 gpurun_out/r02_16_dw3_tests.log
for spec in "64 56 56 96" "64 28 28 192" "64 27 27 384" "64 26 26 768" "64 48 48 96" "64 23 23 384" "16 55 55 384"; do
  set -- $spec
  for op in dwstats dwsplit; do
    for impl in v1 v3; do
      echo -n "$op $impl $spec: "
      CPN_DW_IMPL=$impl timeout 40 python tools/run_op.py $op --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
    done
  done
done 2>&1 | tee gpurun_out/r02_16_dw3_ab.txt
mkdir -p /tmp/ncu_r02
CPN_DW_IMPL=v3 timeout 300 ncu --set full --clock-control none --import-source on -k regex:dwconv7_v3 --launch-skip 2 -c 1 -o /tmp/ncu_r02/dw3c -f python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 3 > gpurun_out/r02_16_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/ncu_r02/dw3c.ncu-rep --page raw --csv > gpurun_out/r02_16_ncu_dw3_raw.csv 2>/dev/null
ncu -i /tmp/ncu_r02/dw3c.ncu-rep --page source --csv > gpurun_out/r02_16_ncu_dw3_source.csv 2>/dev/null
