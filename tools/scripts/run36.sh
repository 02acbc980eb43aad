#!/bin/bash
cd /root/repo
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "ln_im2col2 or mlp_fused" 2>&1 | tail -2
for v in new old; do
  if [ $v = old ]; then export CPN_LNIM_V2=1; else unset CPN_LNIM_V2; fi
  echo "== $v"
  python bench.py 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['kernels'].items() if k in ('ln_im2col2','stem_ln')})"
done
