#!/bin/bash
# Round 2, GPU call 37: state check after the training-head work: full GPU suite, smoke (now with the training head),
# default bench line, reference arm.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q > gpurun_out/r02_37_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/r02_37_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -6
timeout 900 python bench.py > gpurun_out/r02_37_bench_cfg2_n1.json 2> gpurun_out/r02_37_bench_cfg2.err
echo "bench rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_37_bench_cfg2_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"]), "fp32", round((d.get("fp32_mode") or {}).get("value", 0)), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3), "cpu", d.get("cpu_baseline", {}).get("value"))
print("train_head", d.get("train_head"))
PY
