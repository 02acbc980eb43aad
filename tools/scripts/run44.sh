timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "checkpoint or mutation" 2>&1 | tail -8
