export CPN_GEMM_2SM=2
timeout 120 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -8
unset CPN_GEMM_2SM
for m in 0 1; do
echo "2SM=$m"
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 10
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm --M 46656 --N 384 --K 1536 --res 1 --iters 10
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 10
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 3072 --res 1 --iters 10
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm_ln --M 200704 --N 384 --K 96 --iters 10
CPN_GEMM_2SM=$m timeout 60 python tools/run_op.py gemm --M 200704 --N 96 --K 384 --res 1 --iters 10
done
timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 120 python bench.py > gpurun_out/bench_v4.json 2> gpurun_out/bench.err
