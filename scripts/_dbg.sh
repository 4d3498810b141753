PYTHONPATH=. timeout 100 python scripts/_t.py 2>&1 | tail -2
PYTHONPATH=. timeout 100 python scripts/conv_check.py --bench 2>&1 | tail -7
PYTHONPATH=. timeout 200 python scripts/_par.py 2>&1 | tail -2
