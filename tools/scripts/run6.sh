python -m pytest tests/test_kernels_gpu.py -x -q -m gpu 2>&1 | tail -5
python -m pytest tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -5
for nseg in 1 2; do
echo "NSEG=$nseg"
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 28 --W 28 --C 192 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 10
CPN_DWS_NSEG=$nseg python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 10
done
python tools/plan_profile.py > gpurun_out/plan_profile_v3.txt 2> gpurun_out/pp.err
python bench.py > gpurun_out/bench_v3.json 2> gpurun_out/bench.err
head -3 gpurun_out/plan_profile_v3.txt
