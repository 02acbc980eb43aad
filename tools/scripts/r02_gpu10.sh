#!/bin/bash
# Round 2, GPU call 10/11: multi-GPU bench lines (N = $1) for the configs given after it.
cd /root/repo
mkdir -p gpurun_out
N=$1; shift
port=29511
for c in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $N --config $c > gpurun_out/r02_bench_${c}_n${N}.json 2> gpurun_out/r02_bench_${c}_n${N}.err
  echo "$c N=$N rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_bench_${c}_n${N}.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "n_gpus")}, "e2e", d["e2e"]["value"], "e2e_u8", (d.get("e2e_u8_input") or {}).get("value"), "fp32", (d.get("fp32_mode") or {}).get("value"), d["clocks"])
except Exception as e:
    print("no line:", e)
PY
  tail -3 gpurun_out/r02_bench_${c}_n${N}.err
  port=$((port + 1))
done
