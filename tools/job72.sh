set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_cut_sort_adam.py tests/test_gpu_plugin.py -x -q 2>&1 | tail -3 )
timeout 600 python tools/bench_cut.py --k 24 --frames 20 > gpurun_out/cut_k24_s15.json 2> gpurun_out/cut_k24_s15.err; echo "cut exit $?"; tail -n 3 gpurun_out/cut_k24_s15.err; cat gpurun_out/cut_k24_s15.json
timeout 600 python tools/bench_cut.py --k 21 --frames 30 > gpurun_out/cut_k21_s15.json 2> gpurun_out/cut_k21_s15.err; echo "cut exit $?"; cat gpurun_out/cut_k21_s15.json
