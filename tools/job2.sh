set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name 'regex:blend_bwd|blend_fwd|sort_onesweep|preprocess_bwd|preprocess_fwd|build_records|sort_histogram' --launch-skip 36 --launch-count 12 -f -o gpurun_out/r1_full_a python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out
