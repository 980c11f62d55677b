mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/b_plain.json 2> gpurun_out/b_plain.err; echo "plain exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size --clock-control none -c 1500 --csv --log-file gpurun_out/launches_s12.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/ncu_launch.log 2>&1; echo "ncu list exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"blend_fwd|blend_bwd|onesweep|preprocess_bwd|preprocess_fwd|histogram|emit_balanced|scan_sorted|tile_ranges" --launch-skip 42 -c 14 -o gpurun_out/full_s12 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out/*.ncu-rep
