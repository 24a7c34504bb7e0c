set -x
timeout 600 python -m pytest tests/test_box_resize_gpu.py tests/test_script.py -m gpu -q 2>&1 | tail -30
