set -x
mkdir -p gpurun_out/golden
timeout 600 python tests/golden/make_golden.py > gpurun_out/make_golden.log 2>&1; echo "golden exit $?"; tail -8 gpurun_out/make_golden.log
cp gpurun_out/golden/cut_spt.npz gpurun_out/golden/morton.npz tests/golden/
timeout 900 python -m pytest tests/test_gpu_cut_spt.py tests/test_gpu_morton.py tests/test_oracle_golden.py -q -x 2>&1 | tail -15
