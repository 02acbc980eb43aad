timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/v11_pytest_gpu.log
cat gpurun_out/v11_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/v11_bench.json 2> gpurun_out/v11_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/v11_bench.json').read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
