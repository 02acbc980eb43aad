#!/bin/bash
# Round 2, GPU call 6: per-op profiles (all configs), bench lines, ncu captures reduced to CSV on the box (reports stay there).
cd /root/repo
mkdir -p gpurun_out
for c in cfg1 cfg2 cfg3 cfg5; do
  for pr in bf16 fp32; do
    timeout 600 python tools/plan_profile.py --config $c --batch $( [ $c = cfg3 ] && echo 32 || ([ $c = cfg5 ] && echo 16 || echo 64) ) --precision $pr > gpurun_out/r02_6_plan_${c}_${pr}.txt 2>/dev/null; head -1 gpurun_out/r02_6_plan_${c}_${pr}.txt
  done
done
for c in cfg2 cfg1 cfg3 cfg4 cfg5; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_6_bench_$c.json 2> gpurun_out/r02_6_bench_$c.err ; echo "bench $c rc=$?"
done
timeout 900 python bench.py --precision fp32 > gpurun_out/r02_6_bench_cfg2_fp32.json 2> gpurun_out/r02_6_bench_cfg2_fp32.err ; echo "bench fp32 rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_6_bench_cfg2_reference.json 2>/dev/null ; echo "bench reference rc=$?"
timeout 600 python tests/parity_report.py > gpurun_out/r02_6_parity_report.txt 2>/dev/null; echo "parity report rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02_6_ncu_bench.log 2>&1
T=/tmp/ncu_r02; mkdir -p $T
for spec in "96 200704" "192 50176" "384 46656"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none -k regex:mlp_fused --launch-skip 2 -c 1 -o $T/mlp$1 -f python tools/run_op.py mlp --M $2 --C $1 --iters 3 > gpurun_out/r02_6_ncu_mlp$1.log 2>&1
  ncu -i $T/mlp$1.ncu-rep --page raw --csv > gpurun_out/r02_ncu_mlp_fused_$1_raw.csv 2>/dev/null
done
for spec in "96 56" "192 28" "384 27" "768 26"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none -k regex:dwconv7 --launch-skip 2 -c 1 -o $T/dw$1 -f python tools/run_op.py dwstats --B 64 --H $2 --W $2 --C $1 --iters 3 > gpurun_out/r02_6_ncu_dw$1.log 2>&1
  ncu -i $T/dw$1.ncu-rep --page raw --csv > gpurun_out/r02_ncu_dwconv7_stats_$1_raw.csv 2>/dev/null
done
timeout 300 ncu --set full --clock-control none -k regex:conv_gemm --launch-skip 20 -c 1 -o $T/conv -f python tools/plan_profile.py --config cfg3 --batch 32 --precision bf16 --passes 1 > gpurun_out/r02_6_ncu_conv.log 2>&1
ncu -i $T/conv.ncu-rep --page raw --csv > gpurun_out/r02_ncu_conv_gemm_raw.csv 2>/dev/null
du -sh gpurun_out
