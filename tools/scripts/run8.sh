for sm in 0 1; do for mb in 0 32 48 64 96; do
echo "2SM=$sm CHUNK_MB=$mb"
CPN_GEMM_2SM=$sm CPN_L2_CHUNK_MB=$mb timeout 120 python tools/plan_profile.py 2>/dev/null > gpurun_out/pp_${sm}_${mb}.txt; head -1 gpurun_out/pp_${sm}_${mb}.txt
done; done
timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -3
