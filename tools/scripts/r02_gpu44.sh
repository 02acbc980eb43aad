#!/bin/bash
# Round 2, GPU call 44: ncu --set full of the shipped head forward at P = 2048 (one-pass max + sum) and of the loss kernel at
# the ResNet-50 map.
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
cap() {  # name, kernel regex, run_op args...
  local name=$1 rx=$2; shift 2
  timeout 150 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 1 -c 1 -o /tmp/ncu_r02/$name -f python tools/run_op.py "$@" --iters 2 > gpurun_out/r02_44_ncu_$name.log 2>&1
  ncu -i /tmp/ncu_r02/$name.ncu-rep --page raw --csv > gpurun_out/r02_44_ncu_${name}_raw.csv 2>/dev/null
  echo "ncu $name rc=$? $(wc -c < gpurun_out/r02_44_ncu_${name}_raw.csv) bytes"
}
cap headfwd2048 softmax_pool_nchw_kernel head_fwd --B 16 --C 2048 --H 28 --W 28 --N 196
cap loss2048 pipnet_loss_kernel pipnet_loss --B 16 --C 2048 --H 28 --W 28 --N 196
