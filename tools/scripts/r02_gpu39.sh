#!/bin/bash
# Round 2, GPU call 39: second passes of the loss and head-backward kernels walk the channel rows backwards (L2 reuse).
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
timeout 300 python -m pytest tests/test_train_loss_gpu.py -m gpu -q -x > gpurun_out/r02_39_train_tests.log 2>&1
echo "train tests rc=$?"; tail -12 gpurun_out/r02_39_train_tests.log
timeout 300 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_39_train_loss_timing.md 2> gpurun_out/r02_39_timing.err
echo "timing rc=$?"; cat gpurun_out/r02_39_train_loss_timing.md; tail -5 gpurun_out/r02_39_timing.err
cap() {  # name, kernel regex, run_op args...
  local name=$1 rx=$2; shift 2
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 1 -c 1 -o /tmp/ncu_r02/$name -f python tools/run_op.py "$@" --iters 2 > gpurun_out/r02_39_ncu_$name.log 2>&1
  ncu -i /tmp/ncu_r02/$name.ncu-rep --page raw --csv > gpurun_out/r02_39_ncu_${name}_raw.csv 2>/dev/null
  echo "ncu $name rc=$? $(wc -c < gpurun_out/r02_39_ncu_${name}_raw.csv) bytes"
}
cap loss pipnet_loss_kernel pipnet_loss --B 64 --C 768 --H 26 --W 26 --N 200
cap headbwd softmax_pool_nchw_bwd head_bwd --B 64 --C 768 --H 26 --W 26 --N 200
