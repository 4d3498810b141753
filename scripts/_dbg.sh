export PYTHONPATH=.
timeout 100 python scripts/_one_conv.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_tc_kernel -s 2 -c 1 -f -o gpurun_out/r01d_conv_fwd python scripts/_one_conv.py > gpurun_out/ncu_fwd.log 2>&1
tail -3 gpurun_out/ncu_fwd.log
ls -la gpurun_out/*.ncu-rep
