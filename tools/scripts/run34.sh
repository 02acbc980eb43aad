#!/bin/bash
cd /root/repo
for m in 1 2; do
  echo "== CPN_GEMM_2SM=$m"
  CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 30
  CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 30
  CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 3072 --res 1 --iters 30
done
