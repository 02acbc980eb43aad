#!/bin/bash
# Round 2, GPU call 4: v1 depthwise conv with filter-row guards + split output, fused add-on head, all tests, profiles, bench.
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "dwconv7_stats or ln_im2col or ln_gelu_split or addon" > gpurun_out/r02_4_kernels.log 2>&1
echo "kernel tests rc=$?"; tail -6 gpurun_out/r02_4_kernels.log
for shape in "--B 64 --H 27 --W 27 --C 384" "--B 64 --H 56 --W 56 --C 96" "--B 64 --H 28 --W 28 --C 192" "--B 64 --H 26 --W 26 --C 768"; do
  echo "== $shape"
  for n in 0 1 2 3 4; do
    if [ $n = 0 ]; then timeout 120 python tools/run_op.py dwstats $shape --iters 20; else CPN_DWS_NSEG=$n timeout 120 python tools/run_op.py dwstats $shape --iters 20; fi
  done
  timeout 120 python tools/run_op.py dwsplit $shape --iters 20
done 2>&1 | grep -v "^$" | sed 's/{.*}//' | tee gpurun_out/r02_4_dw_ab.txt
timeout 600 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_4_plan_fp32.txt 2>gpurun_out/r02_4_plan_fp32.err; head -1 gpurun_out/r02_4_plan_fp32.txt
timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_4_plan_bf16.txt 2>gpurun_out/r02_4_plan_bf16.err; head -1 gpurun_out/r02_4_plan_bf16.txt
timeout 600 python tools/plan_profile.py --config cfg1 --precision bf16 > gpurun_out/r02_4_plan_cfg1_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_4_plan_cfg1_bf16.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_4_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/r02_4_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02_4_bench_cfg2.json 2> gpurun_out/r02_4_bench_cfg2.err ; echo "bench rc=$?"; cut -c1-200 gpurun_out/r02_4_bench_cfg2.json
for c in cfg1 cfg4; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_4_bench_$c.json 2> gpurun_out/r02_4_bench_$c.err ; echo "bench $c rc=$?"
done
