PYTHONPATH=. timeout 60 python scripts/_conv_abl.py 2>&1 | tail -5
timeout 300 python -m pytest tests/test_conv_gpu.py -x -q 2>&1 | tail -2
