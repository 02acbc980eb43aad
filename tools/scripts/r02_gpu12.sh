#!/bin/bash
# Round 2, GPU call 12: diagnose call 11 (which of v1 / v3 produced no output): short timeouts, full logs, rc and time.
cd /root/repo
mkdir -p gpurun_out
run() {  # label, timeout, command...
  local label=$1 t=$2; shift 2
  local s=$(date +%s.%N)
  timeout $t "$@" > gpurun_out/r02_12_$label.log 2>&1
  local rc=$?
  echo "$label rc=$rc $(echo "$(date +%s.%N) - $s" | bc) s: $(tail -1 gpurun_out/r02_12_$label.log | cut -c1-300)"
}
run v1_a 60 env CPN_DW_IMPL=v1 python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 30
run v3_a 60 env CPN_DW_IMPL=v3 python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 30
run v1_b 60 env CPN_DW_IMPL=v1 python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 30
run v3_b 60 env CPN_DW_IMPL=v3 python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 30
run def_a 60 python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 30
run t1 90 env CPN_DW_IMPL=v3 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "test_dwconv7_stats and 28-28-192"
run t2 90 env CPN_DW_IMPL=v3 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "test_dwconv7_stats and 56-56-96"
