mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/b_plain.json 2> gpurun_out/b_plain.err; echo "plain exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"onesweep|histogram|emit_balanced|preprocess_fwd" --launch-skip 27 -c 9 -o gpurun_out/full_s8 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out/*.ncu-rep
