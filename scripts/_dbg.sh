export PYTHONPATH=.
timeout -k 10 300 python bench.py --only scope_a 2>&1 | tail -2 | cut -c1-1200
