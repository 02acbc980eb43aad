#!/bin/bash
# fused MLP: correctness (both CTA groupings) then timing at the benchmark shapes
cd /root/repo
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "mlp_fused" -s 2>&1 | tail -40
for cg in 2 1; do
  for shape in "46656 384" "50176 192" "200704 96"; do
    set -- $shape
    CPN_MLP_CG=$cg timeout 120 python tools/run_op.py mlp --M $1 --C $2 --iters 20
  done
done
