PYTHONPATH=. timeout 60 python scripts/_conv_abl.py 2>&1 | tail -9
PYTHONPATH=. timeout 100 python scripts/conv_check.py --bench 2>&1 | tail -8
