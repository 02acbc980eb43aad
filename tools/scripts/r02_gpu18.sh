#!/bin/bash
# Round 2, GPU call 18: dwconv7_v3 with a staggered start of the four warps of each scheduler (ns per step in dbg >> 8).
cd /root/repo
mkdir -p gpurun_out
for spec in "64 56 56 96" "64 27 27 384"; do
  set -- $spec
  for ns in 0 300 600 900 1200 1800; do
    echo -n "stagger=$ns $spec: "
    CPN_DW_IMPL=v3 CPN_DW3_DBG=$((ns * 256)) timeout 40 python tools/run_op.py dwstats --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
  done
done 2>&1 | tee gpurun_out/r02_18_dw3_stagger.txt
