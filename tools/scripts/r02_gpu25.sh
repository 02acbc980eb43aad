#!/bin/bash
# Round 2, GPU call 25: fused MLP against the two-GEMM MLP per stage with this round's TMA-epilogue GEMMs (cfg2 plan profile).
cd /root/repo
mkdir -p gpurun_out
CPN_FUSED_MLP=0 timeout 300 python tools/plan_profile.py --precision bf16 > gpurun_out/r02_25_plan_cfg2_bf16_unfused.txt 2>/dev/null
head -1 gpurun_out/r02_25_plan_cfg2_bf16_unfused.txt
grep -n "gemm" gpurun_out/r02_25_plan_cfg2_bf16_unfused.txt | awk '{print $3,$4,$6,$8,$9,$10,$11,$12}' | sort | uniq -c | sort -k2,2 -k3,3n
