set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/v10_pytest_gpu.log
cat gpurun_out/v10_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/v10_bench.json 2> gpurun_out/v10_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/v10_bench_reference.json 2> gpurun_out/v10_bench_reference.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v10_smoke.log 2>&1; tail -4 gpurun_out/v10_smoke.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v10_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/v10_ncu.log 2>&1
python -c "
import json; d=json.loads(open('gpurun_out/v10_bench.json').read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'], d['roofline']['frac'])
d=json.loads(open('gpurun_out/v10_bench_reference.json').read()); print(d['value'], d['cpu_baseline']['cores'])"
