#!/bin/bash
# plain run first (must exit 0), then the launch list of one (eager) step, then full captures of the two conv
# kernels and the coupling adjoint
R=${1:-r02c}
export PYTHONPATH=.
set -x
python bench.py --steps 5 --warmup 3 > gpurun_out/${R}_bench.json 2> gpurun_out/${R}_bench.err || exit 1
python bench.py --only step --steps 1 --warmup 2 > /dev/null 2>&1 || exit 1
CMF_NCU_RANGE=1 timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/${R}_launches_step.csv python bench.py --only step --steps 1 --warmup 2 > gpurun_out/ncu1.log 2>&1
timeout 100 python scripts/one_conv.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_tc_kernel -s 2 -c 1 -f \
    -o gpurun_out/${R}_conv_fwd_tangent python scripts/one_conv.py > gpurun_out/ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_wgrad_tc_kernel -s 2 -c 1 -f \
    -o gpurun_out/${R}_conv_wgrad_tangent python scripts/one_conv.py > gpurun_out/ncu3.log 2>&1
timeout 100 python scripts/one_coupling.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:coupling_bwd_vec4 -s 2 -c 1 -f \
    -o gpurun_out/${R}_coupling_bwd python scripts/one_coupling.py > gpurun_out/ncu4.log 2>&1
timeout 100 python scripts/one_mlp.py || exit 1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mlp_fwd_kernel -s 1 -c 1 -f \
    -o gpurun_out/${R}_mlp_fwd python scripts/one_mlp.py > gpurun_out/ncu5.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mlp_bwd_kernel -s 1 -c 1 -f \
    -o gpurun_out/${R}_mlp_bwd python scripts/one_mlp.py > gpurun_out/ncu6.log 2>&1
timeout 100 python scripts/one_tail.py || exit 1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tail_fwd_kernel -s 2 -c 1 -f \
    -o gpurun_out/${R}_tail_fwd python scripts/one_tail.py > gpurun_out/ncu7.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tail_bwd_tangent -s 2 -c 1 -f \
    -o gpurun_out/${R}_tail_bwd python scripts/one_tail.py > gpurun_out/ncu8.log 2>&1
ls -la gpurun_out/ | tail -12
