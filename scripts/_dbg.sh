export PYTHONPATH=.
timeout -k 10 300 python -m pytest tests/test_conv_gpu.py -x -q -m gpu 2>&1 | tail -3
for d in 0 1; do echo "== CMF_CONV_DBG=$d"; CMF_CONV_DBG=$d timeout -k 5 120 python scripts/conv_check.py --bench 2>&1 | grep "tc  "; done
