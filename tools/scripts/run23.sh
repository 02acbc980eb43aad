#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "mlp_fused" 2>&1 | tail -3
for alt in 1 0; do
  for shape in "46656 384" "50176 192" "200704 96"; do
    set -- $shape
    echo "ALT=$alt"
    CPN_MLP_ALT=$alt timeout 120 python tools/run_op.py mlp --M $1 --C $2 --iters 20
  done
done
CPN_MLP_TS=1 CPN_MLP_DBG=256 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | tail -62
