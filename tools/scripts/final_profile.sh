set -x
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/final_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
timeout 300 python tools/plan_profile.py > gpurun_out/final_plan_profile.txt 2>/dev/null
timeout 300 python tools/plan_profile.py --precision fp32 > gpurun_out/final_plan_profile_fp32.txt 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/final_ncu.log 2>&1
timeout 300 python tests/parity_report.py > gpurun_out/final_parity_report.txt 2>&1
cat gpurun_out/final_pytest_gpu.log; head -c 600 gpurun_out/final_bench.json
