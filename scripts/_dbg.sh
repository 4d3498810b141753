PYTHONPATH=. timeout 100 python scripts/_one_conv.py || exit 1
PYTHONPATH=. timeout 400 ncu --set full --clock-control none --import-source on -k regex:conv3x3_tc_kernel -s 2 -c 1 -o gpurun_out/r01c_conv_fwd_tangent python scripts/_one_conv.py > gpurun_out/ncu4.log 2>&1
PYTHONPATH=. timeout 400 ncu --set full --clock-control none --import-source on -k regex:conv3x3_wgrad_tc_kernel -s 2 -c 1 -o gpurun_out/r01c_conv_wgrad_tangent python scripts/_one_conv.py > gpurun_out/ncu5.log 2>&1
ls -la gpurun_out/*.ncu-rep
