#!/bin/bash
# Round 2, GPU call 11: dwconv7_v3_kernel (lanes over space): kernel tests under CPN_DW_IMPL=v3, then A/B against the
# default (v1) on the benchmark's shapes, both modes.
cd /root/repo
mkdir -p gpurun_out
CPN_DW_IMPL=v3 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats" > gpurun_out/r02_11_dw3_tests.log 2>&1
echo "dw3 tests rc=$?"; tail -25 gpurun_out/r02_11_dw3_tests.log
for spec in "64 56 56 96" "64 28 28 192" "64 27 27 384" "64 26 26 768" "64 48 48 96" "64 24 24 192" "64 23 23 384" "16 112 112 96" "16 55 55 384" "16 54 54 768"; do
  set -- $spec
  for op in dwstats dwsplit; do
    for impl in v1 v3; do
      echo -n "$op $impl $spec: "
      CPN_DW_IMPL=$impl timeout 120 python tools/run_op.py $op --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
    done
  done
done 2>&1 | tee gpurun_out/r02_11_dw3_ab.txt
