#!/bin/bash
cd /root/repo
for rep in 1 2; do
for v in old new_split; do
  case $v in
    old) export CPN_LIB_PATH=/root/repo/ab_tmp/libcpn_old.so; unset CPN_MLP_SPLIT CPN_MLP_ALT;;
    new_split) unset CPN_LIB_PATH CPN_MLP_ALT; export CPN_MLP_SPLIT=1;;
  esac
  echo "== $v"
  timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 30
  timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 30
  timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 30
done
done
