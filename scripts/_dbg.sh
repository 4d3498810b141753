PYTHONPATH=. timeout 100 python scripts/wgrad_check.py 2>&1 | tail -4
for cfg in "3250 28 28 20" "3250 14 14 20" "50 28 28 50"; do PYTHONPATH=. timeout 30 python scripts/_wg2.py $cfg 2>&1 | tail -1; echo "rc=$? ($cfg)"; done
