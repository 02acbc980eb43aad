# ncu --set full captures of the four kernels that make up 90 % of the step, at HEAD (each after its plain run exits 0)
set -x
cap() {  # name, kernel regex, run_op args...
  name=$1; kre=$2; shift 2
  timeout 120 python tools/run_op.py "$@" --iters 3 || return 1
  timeout 400 ncu --set full --clock-control none -k regex:$kre --launch-skip 2 -c 1 -o gpurun_out/v10_$name -f python tools/run_op.py "$@" --iters 3 > gpurun_out/v10_ncu_$name.log 2>&1
  tail -1 gpurun_out/v10_ncu_$name.log
}
cap mlp384 mlp_fused mlp --M 46656 --C 384
cap dws384 dwconv7_stats dwstats --B 64 --H 27 --W 27 --C 384
cap dws768 dwconv7_stats dwstats --B 64 --H 26 --W 26 --C 768
cap fc1_768 gemm_tma_epi gemm_ln --M 43264 --N 3072 --K 768
cap fc2_768 gemm_tma_epi gemm --M 43264 --N 768 --K 3072 --res 1
python tools/ncu_summary.py gpurun_out/v10_mlp384.ncu-rep gpurun_out/v10_dws384.ncu-rep gpurun_out/v10_dws768.ncu-rep gpurun_out/v10_fc1_768.ncu-rep gpurun_out/v10_fc2_768.ncu-rep > gpurun_out/v10_ncu_full_summary.txt 2>&1
for n in dws384 mlp384; do ncu -i gpurun_out/v10_$n.ncu-rep --page details --csv > gpurun_out/v10_${n}_details.csv 2>/dev/null; done
ls -la gpurun_out/*.ncu-rep
rm -f gpurun_out/v10_dws768.ncu-rep gpurun_out/v10_fc1_768.ncu-rep gpurun_out/v10_fc2_768.ncu-rep gpurun_out/v10_mlp384.ncu-rep
cat gpurun_out/v10_ncu_full_summary.txt | head -80
