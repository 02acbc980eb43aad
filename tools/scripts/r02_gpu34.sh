#!/bin/bash
# Round 2, GPU call 34: head forward with the one-pass (online) max + sum for P > 768; bench line with the train_head block.
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
timeout 400 python -m pytest tests/test_train_loss_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_34_train_tests.log 2>&1
echo "train tests rc=$?"; tail -5 gpurun_out/r02_34_train_tests.log
timeout 300 python tools/train_loss_timing.py --iters 20 > gpurun_out/r02_34_train_loss_timing.md 2> gpurun_out/r02_34_timing.err
echo "timing rc=$?"; cat gpurun_out/r02_34_train_loss_timing.md; tail -5 gpurun_out/r02_34_timing.err
timeout 600 python bench.py > gpurun_out/r02_34_bench_cfg2_n1.json 2> gpurun_out/r02_34_bench_cfg2.err
echo "bench rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_34_bench_cfg2_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"]), "fp32", round((d.get("fp32_mode") or {}).get("value", 0)), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
print("train_head", d.get("train_head"))
PY
