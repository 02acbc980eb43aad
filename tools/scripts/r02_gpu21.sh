#!/bin/bash
# Round 2, GPU call 21: ResNet on a bf16 residual stream (bf16 mode): kernel tests, parity tests, cfg3 bench A/B.
cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "conv_bn_act" > gpurun_out/r02_21_conv_tests.log 2>&1
echo "conv tests rc=$?"; tail -8 gpurun_out/r02_21_conv_tests.log
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x > gpurun_out/r02_21_parity.log 2>&1
echo "parity rc=$?"; tail -8 gpurun_out/r02_21_parity.log
timeout 300 python tests/parity_report.py > gpurun_out/r02_21_parity_report.txt 2>&1; grep -i "resnet" gpurun_out/r02_21_parity_report.txt
for v in 1 0; do
  CPN_RESNET_BF16_RES=$v timeout 300 python bench.py --config cfg3 --no-extras > gpurun_out/r02_21_bench_cfg3_res$v.json 2> gpurun_out/r02_21_bench_cfg3_res$v.err
  echo "cfg3 bf16res=$v rc=$?"; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_21_bench_cfg3_res$v.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", d["e2e"]["value"])
PY
done
timeout 300 python tools/plan_profile.py --config cfg3 --precision bf16 > gpurun_out/r02_21_plan_cfg3_bf16.txt 2>/dev/null; head -1 gpurun_out/r02_21_plan_cfg3_bf16.txt
