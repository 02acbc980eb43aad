#!/bin/bash
# Round 2, GPU call 8: fused MLP with 16 epilogue warps at C <= 192: tests + A/B, then full tests and cfg1/cfg2 bench.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "mlp_fused" > gpurun_out/r02_8_mlp_tests.log 2>&1
echo "mlp tests rc=$?"; tail -4 gpurun_out/r02_8_mlp_tests.log
for spec in "96 200704" "192 50176" "96 147456" "192 36864"; do
  set -- $spec
  timeout 120 python tools/run_op.py mlp --M $2 --C $1 --iters 30
  CPN_MLP_EW=8 timeout 120 python tools/run_op.py mlp --M $2 --C $1 --iters 30
done 2>&1 | sed 's/{.*}//' | tee gpurun_out/r02_8_mlp_ab.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_8_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r02_8_pytest_gpu.log
timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_8_plan_cfg2_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_8_plan_cfg2_bf16.txt
timeout 600 python tools/plan_profile.py --config cfg1 --precision bf16 > gpurun_out/r02_8_plan_cfg1_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_8_plan_cfg1_bf16.txt
