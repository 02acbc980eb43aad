python - <<'PY'
import torch
x=torch.empty(1<<30,dtype=torch.uint8,device='cuda')
for name,fn in [("fill (write only)",lambda: x.zero_()),("copy",None)]:
    if fn is None:
        y=torch.empty_like(x); fn=lambda: y.copy_(x)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/10
    print(name, "%.3f ms"%ms, "%.1f GB/s (bytes written%s)"%((1<<30)/ms/1e6*(2 if name=="copy" else 1), " + read" if name=="copy" else ""))
PY
for d in 0 2 4 6; do
echo "DBG=$d"
export CPN_GEMM_DBG=$d
timeout 60 python tools/run_op.py gemm_ln --M 46656 --N 1536 --K 384 --iters 20
timeout 60 python tools/run_op.py gemm_ln --M 200704 --N 384 --K 96 --iters 20
done
