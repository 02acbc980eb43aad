set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/v7_pytest_gpu.log
cat gpurun_out/v7_pytest_gpu.log
CPN_PDL=0 timeout 300 python bench.py > gpurun_out/v7_bench_nopdl.json 2> gpurun_out/v7_bench_nopdl.err
timeout 300 python bench.py > gpurun_out/v7_bench.json 2> gpurun_out/v7_bench.err
CPN_PDL=0 timeout 300 python bench.py > gpurun_out/v7_bench_nopdl2.json 2>/dev/null
timeout 300 python bench.py > gpurun_out/v7_bench2.json 2>/dev/null
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/v7_bench_reference.json 2> gpurun_out/v7_bench_reference.err
timeout 300 python tools/plan_profile.py > gpurun_out/v7_plan_profile.txt 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v7_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/v7_ncu.log 2>&1
timeout 300 python tests/parity_report.py > gpurun_out/v7_parity_report.txt 2>&1
for f in v7_bench_nopdl v7_bench v7_bench_nopdl2 v7_bench2; do python -c "
import json,sys; d=json.load(open('gpurun_out/$f.json')); print('$f', d['value'], d['ms_per_step'], d['e2e']['value'])"; done
