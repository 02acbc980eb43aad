set -x
python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 10
python tools/run_op.py dwstats --B 64 --H 28 --W 28 --C 192 --iters 10
python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 10
python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 10
python tools/run_op.py gemm_ln --M 200704 --N 384 --K 96 --iters 10
python tools/run_op.py gemm --M 200704 --N 384 --K 96 --act 1 --iters 10
python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 10
python tools/run_op.py gemm --M 46656 --N 1536 --K 384 --act 1 --iters 10
python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 10
python tools/run_op.py gemm --M 43264 --N 3072 --K 768 --act 1 --iters 10
ncu --set full --clock-control none --import-source on -k regex:dwconv7_stats -c 1 -s 2 -o gpurun_out/dws384 -f python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 1 > gpurun_out/ncu_dws.log 2>&1
