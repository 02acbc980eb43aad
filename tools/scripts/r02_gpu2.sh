#!/bin/bash
# Round 2, GPU call 2: why is dwconv7_v2 slow? A/B timings + ncu captures; mma.sync microbenchmark; re-test changed kernels.
cd /root/repo
mkdir -p gpurun_out
./tools/micro/mma_bench > gpurun_out/r02_2_mma_bench.txt 2>&1; cat gpurun_out/r02_2_mma_bench.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats or split or ln_im2col" > gpurun_out/r02_2_kernels.log 2>&1
echo "kernel tests rc=$?"; tail -3 gpurun_out/r02_2_kernels.log
for shape in "--B 64 --H 27 --W 27 --C 384" "--B 64 --H 56 --W 56 --C 96" "--B 64 --H 26 --W 26 --C 768"; do
  echo "== $shape"
  python tools/run_op.py dwstats $shape --iters 20
  CPN_DW2_UNROLL=1 python tools/run_op.py dwstats $shape --iters 20
  CPN_DW2_NSEG=1 python tools/run_op.py dwstats $shape --iters 20
  CPN_DW2_NSEG=2 python tools/run_op.py dwstats $shape --iters 20
  CPN_DW_V1=1 python tools/run_op.py dwstats $shape --iters 20
done 2>&1 | tee gpurun_out/r02_2_dw_ab.txt
S3="--B 64 --H 27 --W 27 --C 384 --iters 3"
ncu --set full --clock-control none --import-source on -k regex:dwconv7 --launch-skip 2 -c 1 -o gpurun_out/r02_dw2_roll -f python tools/run_op.py dwstats $S3 > gpurun_out/r02_2_ncu1.log 2>&1
CPN_DW2_UNROLL=1 ncu --set full --clock-control none --import-source on -k regex:dwconv7 --launch-skip 2 -c 1 -o gpurun_out/r02_dw2_unroll -f python tools/run_op.py dwstats $S3 > gpurun_out/r02_2_ncu2.log 2>&1
CPN_DW_V1=1 ncu --set full --clock-control none --import-source on -k regex:dwconv7 --launch-skip 2 -c 1 -o gpurun_out/r02_dw1 -f python tools/run_op.py dwstats $S3 > gpurun_out/r02_2_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
timeout 600 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_2_plan_fp32.txt 2>/dev/null; head -1 gpurun_out/r02_2_plan_fp32.txt
timeout 600 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_2_plan_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_2_plan_bf16.txt
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q > gpurun_out/r02_2_parity.log 2>&1; tail -8 gpurun_out/r02_2_parity.log
