#!/bin/bash
cd /root/repo
for st in 0 40 50 60 -35 -50; do
  echo "== stagger $st"
  export CPN_MLP_STAGGER=$st
  timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 30
  timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 30
  timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 30
done
