for d in 0 1 2 3; do
echo "DBG=$d"
export CPN_GEMM_DBG=$d
timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 20
timeout 60 python tools/run_op.py gemm --M 46656 --N 384 --K 1536 --res 1 --iters 20
done
