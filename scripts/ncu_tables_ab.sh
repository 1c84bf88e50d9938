for v in "NAUTILUS_CSM_TABLE_STORED=1" "NAUTILUS_CSM_YC=32" "NAUTILUS_CSM_YC=16"; do
  env $v ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"pass|normalize|raster" -c 8 --csv --log-file gpurun_out/tab_$v.csv python bench.py --scans 400 --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
done
