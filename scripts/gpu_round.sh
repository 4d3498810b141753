#!/bin/bash
# full GPU suite, then the profile round (bench line, launch list, ncu captures of the two conv kernels)
export PYTHONPATH=.
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -3 gpurun_out/t_all.log
bash scripts/ncu_round.sh r01f 2>&1 | tail -5
timeout -k 10 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01f_bench_reference.json 2>/dev/null; tail -c 600 gpurun_out/r01f_bench_reference.json
