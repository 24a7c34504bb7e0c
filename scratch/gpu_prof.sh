set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python scratch/overhead.py 2>&1 | grep '"use_graph": 2'
