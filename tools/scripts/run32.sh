#!/bin/bash
cd /root/repo
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -2
for v in old new; do
  if [ $v = old ]; then export CPN_LIB_PATH=/root/repo/ab_tmp/libcpn_old.so; else unset CPN_LIB_PATH; fi
  echo "== $v"
  timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 30
  timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 3072 --res 1 --iters 30
  timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 1536 --iters 30
  timeout 60 python tools/run_op.py gemm --M 46656 --N 384 --K 768 --iters 30
  timeout 60 python tools/run_op.py gemm --M 50176 --N 192 --K 384 --iters 30
done
