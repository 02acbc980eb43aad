#!/bin/bash
# Round 2, GPU call 1: kernel tests of the new kernels, parity, bench lines (bf16 / fp32 / v1-depthwise A/B), per-op profiles.
cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/r02_1_gpu.txt 2>&1
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats or split or ln_gelu" > gpurun_out/r02_1_kernels_new.log 2>&1
echo "new-kernel tests rc=$?" ; tail -5 gpurun_out/r02_1_kernels_new.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_1_pytest_gpu.log 2>&1
echo "pytest rc=$?" ; tail -15 gpurun_out/r02_1_pytest_gpu.log
timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_1_plan_bf16.txt 2> gpurun_out/r02_1_plan_bf16.err ; echo "plan bf16 rc=$?"; head -1 gpurun_out/r02_1_plan_bf16.txt
timeout 600 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_1_plan_fp32.txt 2> gpurun_out/r02_1_plan_fp32.err ; echo "plan fp32 rc=$?"; head -1 gpurun_out/r02_1_plan_fp32.txt
CPN_DW_V1=1 timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_1_plan_bf16_dwv1.txt 2>/dev/null ; head -1 gpurun_out/r02_1_plan_bf16_dwv1.txt
CPN_FP32_LEGACY=1 timeout 600 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_1_plan_fp32_legacy.txt 2>/dev/null ; head -1 gpurun_out/r02_1_plan_fp32_legacy.txt
timeout 900 python bench.py > gpurun_out/r02_1_bench_cfg2.json 2> gpurun_out/r02_1_bench_cfg2.err ; echo "bench rc=$?"; cut -c1-400 gpurun_out/r02_1_bench_cfg2.json
for c in cfg1 cfg3 cfg4 cfg5; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_1_bench_$c.json 2> gpurun_out/r02_1_bench_$c.err ; echo "bench $c rc=$?"; cut -c1-300 gpurun_out/r02_1_bench_$c.json
done
