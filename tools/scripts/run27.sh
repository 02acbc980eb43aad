#!/bin/bash
cd /root/repo
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "mlp_fused" 2>&1 | tail -3
for sp in 1 0; do
  echo "SPLIT=$sp"
  CPN_MLP_SPLIT=$sp timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 20
done
timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 20
timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 20
