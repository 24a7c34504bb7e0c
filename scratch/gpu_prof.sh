python -m pytest tests -m gpu -q -x 2>&1 | tail -6
python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k drift -s 2>&1 | grep "E/N"
