#!/bin/bash
# Round 2, GPU call 19: dwconv7_v3 with 256-bit stores: tests, then the phase variants again.
cd /root/repo
mkdir -p gpurun_out
CPN_DW_IMPL=v3 timeout 200 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats" > gpurun_out/r02_19_dw3_tests.log 2>&1
echo "dw3 tests rc=$?"; tail -3 gpurun_out/r02_19_dw3_tests.log
for spec in "64 56 56 96" "64 27 27 384"; do
  set -- $spec
  for dbg in 0 1 2 3; do
    echo -n "dbg=$dbg $spec: "
    CPN_DW_IMPL=v3 CPN_DW3_DBG=$dbg timeout 40 python tools/run_op.py dwstats --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
  done
  echo -n "split $spec: "
  CPN_DW_IMPL=v3 timeout 40 python tools/run_op.py dwsplit --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
done 2>&1 | tee gpurun_out/r02_19_dw3_dbg.txt
