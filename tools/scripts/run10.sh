for ord in 0 1; do for m in 0 1; do
echo "ORDER=$ord 2SM=$m"
export CPN_GEMM_ORDER=$ord CPN_GEMM_2SM=$m
timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 20
timeout 60 python tools/run_op.py gemm --M 46656 --N 384 --K 1536 --res 1 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 43264 --N 3072 --K 768 --iters 20
timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 3072 --res 1 --iters 20
timeout 60 python tools/run_op.py gemm --M 43264 --N 768 --K 1536 --iters 20
done; done
