#!/bin/bash
# ncu evidence for the round: launch list of the bench command + full captures of the fused MLP (stage 3 shape) and
# the dwconv (stage 3 shape). Only after the plain commands have exited 0.
cd /root/repo
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_ncu_launches_v5.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
python tools/run_op.py mlp --M 46656 --C 384 --iters 3 || exit 1
ncu --set full --clock-control none --import-source on -k regex:mlp_fused --launch-skip 2 -c 1 -o gpurun_out/mlp384_v5 -f python tools/run_op.py mlp --M 46656 --C 384 --iters 3 > gpurun_out/ncu_mlp.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mlp_fused --launch-skip 2 -c 1 -o gpurun_out/mlp96_v5 -f python tools/run_op.py mlp --M 200704 --C 96 --iters 3 >> gpurun_out/ncu_mlp.log 2>&1
tail -2 gpurun_out/ncu_mlp.log
