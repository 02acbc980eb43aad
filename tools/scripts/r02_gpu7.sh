#!/bin/bash
# Round 2, GPU call 7: map-writing softmax kernel, stem im2col, latency table, sanitizer, cfg3 / cfg5 profiles.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "softmax_pool or im2col_nchw or conv_bn or addon" > gpurun_out/r02_7_kernels.log 2>&1
echo "kernel tests rc=$?"; tail -4 gpurun_out/r02_7_kernels.log
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q > gpurun_out/r02_7_parity.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r02_7_parity.log
for c in cfg3 cfg5 cfg1; do
  timeout 600 python tools/plan_profile.py --config $c --batch $( [ $c = cfg3 ] && echo 32 || ([ $c = cfg5 ] && echo 16 || echo 64) ) --precision bf16 > gpurun_out/r02_7_plan_${c}_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_7_plan_${c}_bf16.txt
done
timeout 900 python tools/latency_table.py > gpurun_out/r02_7_latency.md 2> gpurun_out/r02_7_latency.err; echo "latency rc=$?"; cat gpurun_out/r02_7_latency.md
# compute-sanitizer on small kernel tests (memcheck: all new kernels; racecheck: the shared-memory / cross-CTA ones)
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "(dwconv7_stats and (12-12-96 or 5-9-192 or 27-27-128 or 6-6-160)) or (gemm_ln_gelu_split and 777) or (conv_bn_act and (1-28-28-64 or 7-8-16)) or (addon_head and 37) or (softmax_pool_and_map and 128) or (mlp_fused and 128-96) or (gemm_split_tma and 900)" > gpurun_out/r02_7_memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/r02_7_memcheck.log | tail -3
timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "(dwconv7_stats and (12-12-96 or 27-27-128)) or (gemm_ln_gelu_split and 777) or (conv_bn_act and 7-8-16) or (addon_head and 37) or (softmax_pool_and_map and 128) or (mlp_fused and 128-96)" > gpurun_out/r02_7_racecheck.log 2>&1
echo "racecheck rc=$?"; grep -E "RACECHECK SUMMARY|passed|failed" gpurun_out/r02_7_racecheck.log | tail -3
du -sh gpurun_out
