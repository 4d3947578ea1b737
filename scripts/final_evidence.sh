#!/bin/bash
# Round-end evidence on one GPU: per-kernel step profile of the final build, then ncu of the convolution kernels on
# five ResNet-50 shapes (plain run first, ncu only if it exited 0).
mkdir -p gpurun_out
timeout 200 python scripts/profile_step.py --rows 45 > gpurun_out/step_profile_final.txt 2>&1
A="28 128 128 3 1 / 56 64 64 3 1 / 14 256 1024 1 1 / 7 512 512 3 1 / 28 256 256 3 2"
timeout 120 python scripts/one_conv_shape.py $A > gpurun_out/one_conv.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --profile-from-start off -k regex:'halo|igemm' -c 40 -f \
  -o gpurun_out/prof_convs python scripts/one_conv_shape.py $A > gpurun_out/ncu_convs.log 2>&1
tail -3 gpurun_out/step_profile_final.txt; cat gpurun_out/one_conv.log; tail -2 gpurun_out/ncu_convs.log
