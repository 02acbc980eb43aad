timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 120 python tools/plan_profile.py > gpurun_out/plan_profile_v4.txt 2> gpurun_out/pp.err
head -4 gpurun_out/plan_profile_v4.txt
timeout 120 python bench.py > gpurun_out/bench_v4.json 2> gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench_v4.json')); print(d['value'], d['ms_per_step'], d['e2e']['value']); [print(k, round(v['ms_per_step'],3)) for k,v in d['kernels'].items()]"
