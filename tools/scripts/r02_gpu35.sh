#!/bin/bash
# Round 2, GPU call 35: loss and head-backward kernels with their tiles in shared memory (cp.async): tests, A/B timing.
cd /root/repo
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_train_loss_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_35_train_tests.log 2>&1
echo "train tests rc=$?"; tail -5 gpurun_out/r02_35_train_tests.log
timeout 300 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_35_train_loss_timing_cache.md 2> gpurun_out/r02_35_timing.err
echo "timing (cache) rc=$?"; cat gpurun_out/r02_35_train_loss_timing_cache.md; tail -5 gpurun_out/r02_35_timing.err
CPN_TRAIN_CACHE=0 timeout 300 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_35_train_loss_timing_nocache.md 2>> gpurun_out/r02_35_timing.err
echo "timing (no cache) rc=$?"; grep -v "^#\|^HBM\|^$" gpurun_out/r02_35_train_loss_timing_nocache.md
