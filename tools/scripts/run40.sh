set -x
nvidia-smi -L | head -8
for n in 8 4 2; do
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
python -c "
import json; d=json.load(open('gpurun_out/scale_n$n.json')); print($n, d['value'], d['ms_per_step'], d['e2e']['value'])"
done
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/scale_ref_n8.json 2> gpurun_out/scale_ref_n8.err
head -c 300 gpurun_out/scale_ref_n8.json
