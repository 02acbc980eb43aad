#!/bin/bash
# Round 2, GPU call 24: GPU input pipeline (cpn_preprocess_u8): tests and throughput at the benchmark shapes.
cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_preprocess_gpu.py -m gpu -q -x > gpurun_out/r02_24_pp_tests.log 2>&1
echo "pp tests rc=$?"; tail -6 gpurun_out/r02_24_pp_tests.log
for spec in "64 375 500 224" "64 224 224 224" "64 256 256 192" "16 768 1024 448"; do
  set -- $spec
  echo -n "preprocess $spec: "
  timeout 60 python tools/run_op.py preprocess --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
done 2>&1 | tee gpurun_out/r02_24_pp_timing.txt
