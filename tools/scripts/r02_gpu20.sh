#!/bin/bash
# Round 2, GPU call 20: dwconv7_v3 as the default where it wins: full GPU suite, bench lines cfg2 / cfg5 / cfg1, plan profiles.
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -m gpu -q > gpurun_out/r02_20_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r02_20_pytest_gpu.log
for c in cfg2 cfg5 cfg1; do
  timeout 600 python bench.py --config $c > gpurun_out/r02_20_bench_${c}_n1.json 2> gpurun_out/r02_20_bench_${c}.err
  echo "$c rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_20_bench_${c}_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"], "fp32", (d.get("fp32_mode") or {}).get("value"), "dw", d["kernels"].get("dwconv7_stats", {}).get("ms_per_step"), d["clocks"]["reasons"])
PY
done
timeout 300 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_20_plan_cfg2_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_20_plan_cfg2_bf16.txt
timeout 300 python tools/plan_profile.py --precision fp32 > gpurun_out/r02_20_plan_cfg2_fp32.txt 2>/dev/null; head -1 gpurun_out/r02_20_plan_cfg2_fp32.txt
