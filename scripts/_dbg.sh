export PYTHONPATH=.
timeout -k 10 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout -k 10 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/b.json 2>/dev/null; python -c "
import json; d=json.loads([l for l in open('gpurun_out/b.json') if l.startswith('{')][-1]); print(d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],1) for k,v in d['conv_kernels'].items()})"
