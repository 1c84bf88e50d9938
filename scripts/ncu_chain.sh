#!/bin/bash
# ncu full captures of the matching -> residual chain kernels (K7, K7b, K10 resident) over a
# 2000-pose trajectory; one launch per kernel.  Run under gpurun; reports land in gpurun_out/.
CMD="python scripts/bench_trajectory.py --poses 500"
$CMD > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"nn_match_kernel" --launch-skip 9 --launch-count 1 -f -o gpurun_out/r1s_chain_nn $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"corr_count_kernel|corr_write_kernel|residual_resident_kernel" --launch-count 3 -f -o gpurun_out/r1s_chain_res $CMD > /dev/null 2>&1
ls -la gpurun_out/r1s_chain_*.ncu-rep
