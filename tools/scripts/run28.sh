#!/bin/bash
cd /root/repo
for sp in 1 0; do
  echo "SPLIT=$sp"
  CPN_MLP_SPLIT=$sp CPN_MLP_TS=1 CPN_MLP_DBG=512 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | grep -A4 "tiles of"
  CPN_MLP_SPLIT=$sp timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 20
done
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv
