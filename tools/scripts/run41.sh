set -x
timeout 300 python bench.py > gpurun_out/v9_bench.json 2> gpurun_out/v9_bench.err
wc -l gpurun_out/v9_bench.json
python -c "
import json; d=json.loads(open('gpurun_out/v9_bench.json').read()); print(1, d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/v9_scale_n2.json 2> gpurun_out/v9_scale_n2.err
wc -l gpurun_out/v9_scale_n2.json
python -c "
import json; d=json.loads(open('gpurun_out/v9_scale_n2.json').read()); print(2, d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
timeout 300 python -m pytest tests/test_eval_metrics.py -x -q -m gpu 2>&1 | tail -5
timeout 120 python - > gpurun_out/v9_eval_timing.txt 2>&1 <<'PY'
import torch, time
from count_pipnet_b200 import eval_metrics, ops
from oracle import reference_eval
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(5)
B, P, K = 64, 768, 200
W = torch.where(torch.rand(K, P, generator=g) < 0.01, torch.rand(K, P, generator=g) * 10, torch.zeros(()))
for dens in (0.2, 0.02):
    pooled = torch.where(torch.rand(B, P, generator=g) < 1 - dens, torch.zeros(()), torch.rand(B, P, generator=g))
    out = pooled @ W.t(); ys = torch.randint(0, K, (B,), generator=g)
    ev = eval_metrics.BatchEvaluator(K, P, dev)
    a = [t.to(dev) for t in (pooled, out, W, ys)]
    for _ in range(5): ev.update(*a, 2.0)
    torch.cuda.synchronize()
    res = ev._buffers(B)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): ops.eval_batch(a[0], a[1], a[2], a[3], 2.0, 1e-3, res, cm=ev.cm)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    t0 = time.perf_counter()
    for _ in range(3): reference_eval.batch_statistics(pooled, out, W, ys, 2.0)
    cpu_ms = (time.perf_counter() - t0) / 3 * 1e3
    print(f"cpn_eval_batch B={B} P={P} K={K}, {dens:.0%} of pooled non-zero: {ms*1e3:.1f} us per batch (memset + 2 kernels; {4.0*B*K*P/ms/1e6:.0f} GB/s of the dense [B,K,P] weight reads, table L2-resident); reference formulation on {torch.get_num_threads()} host threads: {cpu_ms:.1f} ms")
PY
cat gpurun_out/v9_eval_timing.txt
