timeout 200 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -2
for ew in 8 16; do
echo "EW=$ew"
export CPN_GEMM_EW=$ew
timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 200704 --N 384 --K 96 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 50176 --N 768 --K 192 --iters 20
done
