timeout 600 python -m pytest tests/test_pair_evaluators_gpu.py tests/test_script.py -m gpu -q 2>&1 | tail -6
