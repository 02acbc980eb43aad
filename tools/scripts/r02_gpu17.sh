#!/bin/bash
# Round 2, GPU call 17: where dwconv7_v3 loses time: variants without the output phase / the FMAs / the data wait.
cd /root/repo
mkdir -p gpurun_out
for spec in "64 56 56 96" "64 27 27 384"; do
  set -- $spec
  for dbg in 0 1 2 4 5 3; do
    echo -n "dbg=$dbg $spec: "
    CPN_DW_IMPL=v3 CPN_DW3_DBG=$dbg timeout 40 python tools/run_op.py dwstats --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
  done
done 2>&1 | tee gpurun_out/r02_17_dw3_dbg.txt
