for pf in 0 2 4 6; do for nseg in 1 2 3; do
echo "PF=$pf NSEG=$nseg"
CPN_DWS_PF=$pf CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 10
done; done
for pf in 0 4; do for nseg in 1 2 4; do
echo "PF=$pf NSEG=$nseg"
CPN_DWS_PF=$pf CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 10
CPN_DWS_PF=$pf CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 10
done; done
python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "dwconv7_stats" 2>&1 | tail -3
