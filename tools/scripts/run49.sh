for n in 4 2; do
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/v11_scale_n$n.json 2> gpurun_out/v11_scale_n$n.err
wc -l gpurun_out/v11_scale_n$n.json
python -c "
import json; d=json.loads(open('gpurun_out/v11_scale_n$n.json').read()); print($n, d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
done
