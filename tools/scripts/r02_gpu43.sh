#!/bin/bash
# Round 2, GPU call 43: head backward with three channel rows in flight per thread and tensor: tests, timing.
cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_train_loss_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_43_train_tests.log 2>&1
echo "train tests rc=$?"; tail -4 gpurun_out/r02_43_train_tests.log
timeout 200 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_43_train_loss_timing.md 2> gpurun_out/r02_43_timing.err
echo "timing rc=$?"; grep "backward\|loss +" gpurun_out/r02_43_train_loss_timing.md | cut -d'|' -f2-5
