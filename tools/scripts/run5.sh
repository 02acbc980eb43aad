python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "dwconv7_stats" 2>&1 | tail -3
for nseg in 1 2 3; do
echo "NSEG=$nseg"
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 28 --W 28 --C 192 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 10
done
ncu --set full --clock-control none --import-source on -k regex:dwconv7_stats -c 1 -s 2 -o gpurun_out/dws384c -f python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 1 > gpurun_out/ncu_dws.log 2>&1
