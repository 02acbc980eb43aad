#!/bin/bash
cd /root/repo
for dbg in 0 8 16 24; do
  echo "DBG=$dbg"
  CPN_MLP_DBG=$dbg timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 20
  CPN_MLP_TS=1 CPN_MLP_DBG=$((dbg+256)) timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | grep -A3 "tiles of"
done
