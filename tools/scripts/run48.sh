for ns in 1 2 3 4; do
export CPN_DWS_NSEG=$ns
echo "NSEG=$ns"
timeout 60 python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 30 | grep -o ": [0-9.]* ms"
timeout 60 python tools/run_op.py dwstats --B 64 --H 28 --W 28 --C 192 --iters 30 | grep -o ": [0-9.]* ms"
timeout 60 python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 30 | grep -o ": [0-9.]* ms"
timeout 60 python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 30 | grep -o ": [0-9.]* ms"
done
