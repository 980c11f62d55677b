mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none --launch-skip 120 -c 60 --csv --log-file gpurun_out/launches2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu exit $?"
