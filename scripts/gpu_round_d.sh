#!/bin/bash
# round r01d: packed-stream conv forward + ragged sub-chunk skip in the wgrad.  Kernel tests first
# (own timeouts: a protocol bug would hang, not fail), then timings, the full GPU suite and the bench.
export PYTHONPATH=.
mkdir -p gpurun_out
timeout -k 10 400 python -m pytest tests/test_conv_gpu.py -x -q -m gpu > gpurun_out/t_conv.log 2>&1
rc=$?; echo "conv tests rc=$rc"; tail -5 gpurun_out/t_conv.log
[ $rc -ne 0 ] && exit 1
timeout -k 10 300 python scripts/conv_check.py --bench > gpurun_out/conv_check.log 2>&1; tail -8 gpurun_out/conv_check.log
timeout -k 10 300 python scripts/wgrad_check.py --bench > gpurun_out/wgrad_check.log 2>&1; tail -8 gpurun_out/wgrad_check.log
timeout -k 10 1200 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -5 gpurun_out/t_all.log
timeout -k 10 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r01d_bench.json 2> gpurun_out/r01d_bench.err; echo "bench rc=$?"
cat gpurun_out/r01d_bench.json | head -c 1500
