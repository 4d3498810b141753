export PYTHONPATH=.
timeout -k 10 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "ood" 2>&1 | tail -15
