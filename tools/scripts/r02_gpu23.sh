#!/bin/bash
# Round 2, GPU call 23: dwconv7_v3 bf16 mode with the one-round bf16 scratch; PDL A/B under CUDA graphs.
cd /root/repo
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv7_stats" > gpurun_out/r02_23_dw3_tests.log 2>&1
echo "dw tests rc=$?"; tail -3 gpurun_out/r02_23_dw3_tests.log
for spec in "64 56 56 96" "64 28 28 192" "64 27 27 384" "64 26 26 768"; do
  set -- $spec
  echo -n "dwstats v3 $spec: "
  CPN_DW_IMPL=v3 timeout 40 python tools/run_op.py dwstats --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
done 2>&1 | tee gpurun_out/r02_23_dw3.txt
for c in cfg2 cfg1; do
  for pdl in 0 1; do
    CPN_PDL=$pdl timeout 300 python bench.py --config $c --no-extras > gpurun_out/r02_23_bench_${c}_pdl$pdl.json 2> gpurun_out/r02_23_bench_${c}_pdl$pdl.err
    echo -n "$c pdl=$pdl rc=$?: "; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_23_bench_${c}_pdl$pdl.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"], "separated", d.get("step_ms_event_separated"))
PY
  done
done
