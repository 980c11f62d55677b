set -x
mkdir -p gpurun_out
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name 'regex:blend_bwd|blend_fwd|preprocess_fwd|build_records|sort_onesweep|emit_tiles|scan_sorted' --launch-skip 33 --launch-count 11 -f -o gpurun_out/r1_full_b python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out
