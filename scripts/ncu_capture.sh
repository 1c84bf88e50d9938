#!/bin/bash
# ncu evidence for one build (run under gpurun, ONE GPU): launch list + one --set full
# capture per step kernel of `python bench.py --scans N` (N tables, 5 N pairs per
# step; default N = 2000, the bench size).  Reports land in gpurun_out/<tag>_*; summarise here with
#   python scripts/ncu_summary.py <tag> gpurun_out/<tag>_launches.csv gpurun_out/<tag>_full.ncu-rep [bench.json] [N]
TAG=${1:-r2}
SCANS=${2:-2000}   # full bench size: waves of >= 1200 tables take the two-run table path
CMD="python bench.py --scans $SCANS --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_ncu_plain.json 2> gpurun_out/${TAG}_ncu_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
# one launch of each kernel from the second (timed) step: skip the warm-up step's launches
ncu --set full --clock-control none --import-source on \
    -k regex:"raster_kernel|xpass_kernel|ypass2_kernel|low_finalize_kernel|low_score_kernel|refine_first_kernel|refine_spec_kernel|refine_finish_kernel" \
    --launch-skip 9 --launch-count 9 -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/${TAG}_*
