timeout 300 python -m pytest tests/test_eval_metrics.py tests/test_search.py -x -q -m gpu 2>&1 | tail -8
