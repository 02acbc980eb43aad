#!/bin/bash
# Round 2, GPU call 40: final state: full GPU suite, smoke, default bench line, reference arm, ncu launch list of smoke.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q > gpurun_out/r02_40_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/r02_40_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -5
timeout 900 python bench.py > gpurun_out/r02_40_bench_cfg2_n1.json 2> gpurun_out/r02_40_bench_cfg2.err
echo "bench rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_40_bench_cfg2_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"]), "fp32", round((d.get("fp32_mode") or {}).get("value", 0)), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
print("train_head", d.get("train_head"))
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_40_bench_reference.json 2> gpurun_out/r02_40_bench_reference.err
echo "reference arm rc=$?"; tail -c 400 gpurun_out/r02_40_bench_reference.json
timeout 300 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_40_train_loss_timing.md 2>/dev/null; grep "cfg2\|cfg3" -A2 gpurun_out/r02_40_train_loss_timing.md | head -8
