export PYTHONPATH=.
timeout -k 10 300 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -5
timeout -k 10 300 python bench.py --only scope_a 2>&1 | tail -1 | cut -c1-700
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'chol|gram' -c 12 --csv --log-file gpurun_out/scope_a_small.csv python bench.py --only scope_a > /dev/null 2>&1
python - <<'P'
import csv
rows=list(csv.reader(l for l in open('gpurun_out/scope_a_small.csv') if l.startswith('"')))
h=rows[0]; ci={x:i for i,x in enumerate(h)}
for r in rows[1:]:
    print(r[ci['Kernel Name']][:40], r[ci['Metric Value']], r[ci['Metric Unit']], r[ci['Grid Size']], r[ci['Block Size']])
P
