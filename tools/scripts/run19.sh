#!/bin/bash
# fused MLP: switch-off experiments + one ncu capture
cd /root/repo
for dbg in 0 1 2 3 4 6 7; do
  echo "DBG=$dbg"
  CPN_MLP_DBG=$dbg timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 20
  CPN_MLP_DBG=$dbg CPN_MLP_CG=1 timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 20
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_fused -c 1 -o gpurun_out/mlp384 -f python tools/run_op.py mlp --M 46656 --C 384 --iters 1 > gpurun_out/ncu_mlp.log 2>&1
CPN_MLP_CG=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_fused -c 1 -o gpurun_out/mlp192 -f python tools/run_op.py mlp --M 50176 --C 192 --iters 1 >> gpurun_out/ncu_mlp.log 2>&1
tail -3 gpurun_out/ncu_mlp.log
