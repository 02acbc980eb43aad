#!/bin/bash
# Round 2, GPU call 5: state after the depthwise revert: tests, per-op profiles (all configs), bench lines, ncu captures.
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_5_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r02_5_pytest_gpu.log
for c in cfg1 cfg2 cfg3 cfg5; do
  for pr in bf16 fp32; do
    timeout 600 python tools/plan_profile.py --config $c --batch $( [ $c = cfg3 ] && echo 32 || ([ $c = cfg5 ] && echo 16 || echo 64) ) --precision $pr > gpurun_out/r02_5_plan_${c}_${pr}.txt 2>/dev/null; head -1 gpurun_out/r02_5_plan_${c}_${pr}.txt
  done
done
for c in cfg2 cfg1 cfg3 cfg4 cfg5; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_5_bench_$c.json 2> gpurun_out/r02_5_bench_$c.err ; echo "bench $c rc=$?"
done
timeout 900 python bench.py --precision fp32 > gpurun_out/r02_5_bench_cfg2_fp32.json 2> gpurun_out/r02_5_bench_cfg2_fp32.err ; echo "bench fp32 rc=$?"
# ncu: launch list of the bench command, then full captures of the fused MLP at the three widths and of the depthwise conv
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02_5_ncu_bench.log 2>&1
for spec in "96 200704" "192 50176" "384 46656"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:mlp_fused --launch-skip 2 -c 1 -o gpurun_out/r02_mlp$1 -f python tools/run_op.py mlp --M $2 --C $1 --iters 3 > gpurun_out/r02_5_ncu_mlp$1.log 2>&1
done
for spec in "96 56" "192 28" "768 26"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none -k regex:dwconv7 --launch-skip 2 -c 1 -o gpurun_out/r02_dw1_c$1 -f python tools/run_op.py dwstats --B 64 --H $2 --W $2 --C $1 --iters 3 > gpurun_out/r02_5_ncu_dw$1.log 2>&1
done
ls -la gpurun_out/*.ncu-rep | tail -8
