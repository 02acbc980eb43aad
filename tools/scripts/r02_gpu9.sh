#!/bin/bash
# Round 2, GPU call 9: the allocation-free softmax+pool entry point (alone in a fresh process, then at the end of the
# full GPU suite run with -x as the driver does), smoke() and the default bench line at HEAD.
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_9_no_alloc_alone.log 2>&1
echo "no_alloc alone rc=$?"; tail -15 gpurun_out/r02_9_no_alloc_alone.log
timeout 1800 python -m pytest tests -x -m gpu -q > gpurun_out/r02_9_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/r02_9_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02_9_bench_cfg2_n1.json 2> gpurun_out/r02_9_bench_cfg2_n1.err
echo "bench rc=$?"; cat gpurun_out/r02_9_bench_cfg2_n1.json
