#!/bin/bash
# Round 2, GPU call 27: input pipeline with the per-call table: tests, timing, cfg1 / cfg2 bench lines, two ncu captures.
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
timeout 300 python -m pytest tests/test_preprocess_gpu.py -m gpu -q -x > gpurun_out/r02_27_pp_tests.log 2>&1
echo "pp tests rc=$?"; tail -3 gpurun_out/r02_27_pp_tests.log
for spec in "64 375 500 224" "64 224 224 224" "64 224 224 192" "16 512 512 448"; do
  set -- $spec
  echo -n "preprocess $spec: "
  timeout 60 python tools/run_op.py preprocess --B $1 --H $2 --W $3 --C $4 --iters 30 2>&1 | tail -1 | sed 's/{.*}//'
done 2>&1 | tee gpurun_out/r02_27_pp_timing.txt
for c in cfg1 cfg2; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_27_bench_${c}_n1.json 2> gpurun_out/r02_27_bench_$c.err
  echo -n "bench $c rc=$?: "; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_27_bench_${c}_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"]), "e2e_u8", round((d.get("e2e_u8_input") or {}).get("value", 0)), "fp32", round((d.get("fp32_mode") or {}).get("value", 0)))
PY
done
cap() {
  local name=$1 rx=$2; shift 2
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 2 -c 1 -o /tmp/ncu_r02/$name -f python tools/run_op.py "$@" --iters 3 > gpurun_out/r02_27_ncu_$name.log 2>&1
  ncu -i /tmp/ncu_r02/$name.ncu-rep --page raw --csv > gpurun_out/r02_26_ncu_${name}_raw.csv 2>/dev/null
  echo "ncu $name rc=$? $(wc -c < gpurun_out/r02_26_ncu_${name}_raw.csv) bytes"
}
cap dw3_768 dwconv7_v3 dwstats --B 64 --H 26 --W 26 --C 768
cap pp pp_resize preprocess --B 64 --H 375 --W 500 --C 224
