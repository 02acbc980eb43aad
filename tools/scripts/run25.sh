#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "mlp_fused" 2>&1 | tail -2
for shape in "46656 384" "50176 192" "200704 96"; do
  set -- $shape
  timeout 120 python tools/run_op.py mlp --M $1 --C $2 --iters 20
done
CPN_MLP_TS=1 CPN_MLP_DBG=256 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | grep -A4 "tiles of"
python bench.py 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})"
