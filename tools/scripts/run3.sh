set -x
python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "dwconv7_stats or gemm_ln_gelu or gelu" 2>&1 | tail -5
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
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py > gpurun_out/bench_v3.json 2> gpurun_out/bench.err
ncu --set full --clock-control none --import-source on -k regex:dwconv7_stats -c 1 -s 2 -o gpurun_out/dws384b -f python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 1 > gpurun_out/ncu_dws.log 2>&1
