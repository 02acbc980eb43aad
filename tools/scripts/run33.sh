#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "mlp" -s 2>&1 | grep -E "vs_two|passed|failed" | head -8
for v in old new; do
  if [ $v = old ]; then export CPN_LIB_PATH=/root/repo/ab_tmp/libcpn_old.so; else unset CPN_LIB_PATH; fi
  echo "== $v"
  timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 30
  timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 30
  timeout 120 python tools/run_op.py mlp --M 200704 --C 96 --iters 30
done
unset CPN_LIB_PATH
timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python tests/parity_report.py > gpurun_out/parity_report_v5.txt 2>&1; grep bf16 gpurun_out/parity_report_v5.txt | head -20
