set -x
timeout 600 python -m pytest tests/test_eval_metrics.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/v8_pytest_eval.log
cat gpurun_out/v8_pytest_eval.log
timeout 120 python - > gpurun_out/v8_eval_timing.txt 2>&1 <<'PY'
import torch, time
from count_pipnet_b200 import eval_metrics
from oracle import reference_eval
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(5)
B, P, K = 64, 768, 200
W = torch.where(torch.rand(K, P, generator=g) < 0.01, torch.rand(K, P, generator=g) * 10, torch.zeros(()))
pooled = torch.where(torch.rand(B, P, generator=g) < 0.8, torch.zeros(()), torch.rand(B, P, generator=g))
out = pooled @ W.t(); ys = torch.randint(0, K, (B,), generator=g)
ev = eval_metrics.BatchEvaluator(K, P, dev)
a = [t.to(dev) for t in (pooled, out, W, ys)]
for _ in range(5): ev.update(*a, 2.0)
torch.cuda.synchronize()
from count_pipnet_b200 import ops
res = ev._buffers(B)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): ops.eval_batch(a[0], a[1], a[2], a[3], 2.0, 1e-3, res, cm=ev.cm)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
print(f"cpn_eval_batch B={B} P={P} K={K}: {ms*1e3:.1f} us per batch (memset + 2 kernels), {4.0*B*K*P/ms/1e6:.1f} GB/s of [B,K,P] weight reads (L2-resident table)")
t0 = time.perf_counter()
for _ in range(3): reference_eval.batch_statistics(pooled, out, W, ys, 2.0)
print(f"oracle (reference formulation, [K,B,P] tensor, CPU {torch.get_num_threads()} threads): {(time.perf_counter()-t0)/3*1e3:.1f} ms per batch")
PY
cat gpurun_out/v8_eval_timing.txt
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/v8_pytest_gpu.log
cat gpurun_out/v8_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/v8_bench.json 2> gpurun_out/v8_bench.err
python -c "
import json; d=json.load(open('gpurun_out/v8_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v8_smoke.log 2>&1; tail -3 gpurun_out/v8_smoke.log
