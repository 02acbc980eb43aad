for tw in 4 3; do
export CPN_DWS_TW=$tw
echo "TW=$tw"
timeout 200 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "dwconv7_stats" 2>&1 | tail -2
timeout 60 python tools/run_op.py dwstats --B 64 --H 56 --W 56 --C 96 --iters 20
timeout 60 python tools/run_op.py dwstats --B 64 --H 28 --W 28 --C 192 --iters 20
timeout 60 python tools/run_op.py dwstats --B 64 --H 27 --W 27 --C 384 --iters 20
timeout 60 python tools/run_op.py dwstats --B 64 --H 26 --W 26 --C 768 --iters 20
timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('bench', d['value'], d['ms_per_step'], d['kernels']['dwconv7_stats']['ms_per_step'])"
done
