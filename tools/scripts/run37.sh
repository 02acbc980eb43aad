#!/bin/bash
cd /root/repo
for v in old new old new; do
  if [ $v = old ]; then export CPN_LNIM_V2=1; else unset CPN_LNIM_V2; fi
  echo "== $v"
  python bench.py --steps 40 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['kernels'].items() if k in ('ln_im2col2','stem_ln','mlp_fused','dwconv7_stats')})"
done
