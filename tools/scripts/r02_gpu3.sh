#!/bin/bash
# Round 2, GPU call 3: tensor-core depthwise conv (dwconv_mma.cuh): tests, A/B timings against v1 / v2, step profiles.
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats or ln_im2col or ln_gelu_split or addon" > gpurun_out/r02_3_kernels.log 2>&1
echo "kernel tests rc=$?"; tail -12 gpurun_out/r02_3_kernels.log
for shape in "--B 64 --H 27 --W 27 --C 384" "--B 64 --H 56 --W 56 --C 96" "--B 64 --H 28 --W 28 --C 192" "--B 64 --H 26 --W 26 --C 768" "--B 16 --H 54 --W 54 --C 768" "--B 1 --H 27 --W 27 --C 384"; do
  echo "== $shape"
  timeout 120 python tools/run_op.py dwstats $shape --iters 20
  CPN_DWM_NSEG=1 timeout 120 python tools/run_op.py dwstats $shape --iters 20
  CPN_DW_IMPL=v1 timeout 120 python tools/run_op.py dwstats $shape --iters 20
  timeout 120 python tools/run_op.py dwsplit $shape --iters 20
done 2>&1 | grep -v "^$" | sed 's/{.*}//' | tee gpurun_out/r02_3_dw_ab.txt
timeout 600 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_3_plan_fp32.txt 2>gpurun_out/r02_3_plan_fp32.err; head -1 gpurun_out/r02_3_plan_fp32.txt
timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_3_plan_bf16.txt 2>gpurun_out/r02_3_plan_bf16.err; head -1 gpurun_out/r02_3_plan_bf16.txt
S3="--B 64 --H 27 --W 27 --C 384 --iters 3"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dwconv7 --launch-skip 2 -c 1 -o gpurun_out/r02_dwm_s3 -f python tools/run_op.py dwstats $S3 > gpurun_out/r02_3_ncu1.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_3_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/r02_3_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02_3_bench_cfg2.json 2> gpurun_out/r02_3_bench_cfg2.err ; echo "bench rc=$?"; cut -c1-200 gpurun_out/r02_3_bench_cfg2.json
