#!/bin/bash
# Round 2, GPU call 42: A/B of the CTA size of the loss and head-backward kernels (512 / 256 / 128 threads).
cd /root/repo
mkdir -p gpurun_out
for t in 512 256 128; do
  CPN_TRAIN_THREADS=$t timeout 200 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_42_timing_t$t.md 2> gpurun_out/r02_42_err_t$t.txt
  echo "== threads $t rc=$?"; grep "loss +\|backward" gpurun_out/r02_42_timing_t$t.md | cut -d'|' -f2-5
done
