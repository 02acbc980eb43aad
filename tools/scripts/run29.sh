#!/bin/bash
cd /root/repo; timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "mlp_fused" 2>&1 | tail -2
cd /root/repo
for rep in 1; do
for v in old new_split new_nosplit new_noalt; do
  case $v in
    old) export CPN_LIB_PATH=/root/repo/ab_tmp/libcpn_old.so; unset CPN_MLP_SPLIT CPN_MLP_ALT;;
    new_split) unset CPN_LIB_PATH CPN_MLP_ALT; export CPN_MLP_SPLIT=1;;
    new_nosplit) unset CPN_LIB_PATH CPN_MLP_ALT; export CPN_MLP_SPLIT=0;;
    new_noalt) unset CPN_LIB_PATH; export CPN_MLP_SPLIT=0 CPN_MLP_ALT=0;;
  esac
  echo "== $v"
  timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 30
  timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 30
  timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 30
done
done
unset CPN_LIB_PATH CPN_MLP_ALT; export CPN_MLP_SPLIT=1
CPN_MLP_TS=1 CPN_MLP_DBG=512 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 > gpurun_out/trace_split.log 2>&1
grep -A4 "tiles of" gpurun_out/trace_split.log
