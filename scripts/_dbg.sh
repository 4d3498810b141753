export PYTHONPATH=.
timeout -k 10 900 python scripts/bench_cifar.py 2> /dev/null | tee gpurun_out/cifar_sweep.jsonl
