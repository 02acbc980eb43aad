#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_kernels_gpu.py -x -q -m gpu -k "not mlp_fused" 2>&1 | tail -2
for f in 2 3 4 6; do
  echo "== pool chunk factor $f"
  CPN_POOL_CHUNKS=$f python bench.py 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['kernels'].items() if k in ('softmax_pool','head_tail')})"
done
