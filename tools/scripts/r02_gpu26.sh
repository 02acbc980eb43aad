#!/bin/bash
# Round 2, GPU call 26: final state: full GPU suite, smoke, bench lines (all configs + fp32), plan profiles, parity report,
# latency table, ncu launch list and --set full captures of the kernels that changed this round.
cd /root/repo
mkdir -p gpurun_out /tmp/ncu_r02
timeout 1500 python -m pytest tests -x -m gpu -q > gpurun_out/r02_26_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/r02_26_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
for c in cfg2 cfg1 cfg3 cfg4 cfg5; do
  timeout 900 python bench.py --config $c > gpurun_out/r02_26_bench_${c}_n1.json 2> gpurun_out/r02_26_bench_$c.err
  echo -n "bench $c rc=$?: "; python - <<PY
import json
d = json.loads(open("gpurun_out/r02_26_bench_${c}_n1.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"]), "e2e_u8", round((d.get("e2e_u8_input") or {}).get("value", 0)), "fp32", round((d.get("fp32_mode") or {}).get("value", 0)), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3), d["clocks"]["reasons"])
PY
done
timeout 900 python bench.py --precision fp32 > gpurun_out/r02_26_bench_cfg2_fp32_n1.json 2> gpurun_out/r02_26_bench_cfg2_fp32.err; echo "bench fp32 rc=$?"
for c in cfg1 cfg2 cfg3 cfg5; do
  for pr in bf16 fp32; do
    timeout 600 python tools/plan_profile.py --config $c --batch $( [ $c = cfg3 ] && echo 32 || ([ $c = cfg5 ] && echo 16 || echo 64) ) --precision $pr > gpurun_out/r02_26_plan_${c}_${pr}.txt 2>/dev/null; head -1 gpurun_out/r02_26_plan_${c}_${pr}.txt
  done
done
timeout 600 python tests/parity_report.py > gpurun_out/r02_26_parity_report.txt 2>&1; grep -c PASS gpurun_out/r02_26_parity_report.txt; grep -c FAIL gpurun_out/r02_26_parity_report.txt
timeout 600 python tools/latency_table.py > gpurun_out/r02_26_latency.md 2>/dev/null; tail -8 gpurun_out/r02_26_latency.md
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_26_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02_26_ncu_bench.log 2>&1
echo "ncu launches rc=$?"
cap() {  # name, kernel regex, run_op args...
  local name=$1 rx=$2; shift 2
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$rx --launch-skip 2 -c 1 -o /tmp/ncu_r02/$name -f python tools/run_op.py "$@" --iters 3 > gpurun_out/r02_26_ncu_$name.log 2>&1
  ncu -i /tmp/ncu_r02/$name.ncu-rep --page raw --csv > gpurun_out/r02_26_ncu_${name}_raw.csv 2>/dev/null
  echo "ncu $name rc=$? $(wc -c < gpurun_out/r02_26_ncu_${name}_raw.csv) bytes"
}
cap dw3_96 dwconv7_v3 dwstats --B 64 --H 56 --W 56 --C 96
cap dw3_192 dwconv7_v3 dwstats --B 64 --H 28 --W 28 --C 192
cap dw3_384 dwconv7_v3 dwstats --B 64 --H 27 --W 27 --C 384
cap dw3_768 dwconv7_v3 dwstats --B 64 --H 26 --W 26 --C 768
cap dw3s_384 dwconv7_v3 dwsplit --B 64 --H 27 --W 27 --C 384
cap mlp384 mlp_fused mlp --M 46656 --C 384
cap pp pp_resize preprocess --B 64 --H 375 --W 500 --C 224
