set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python scratch/overhead.py 2>&1 | tail -8
python scratch/sweep.py lj64k 2000 2>&1 | tail -6
