mkdir -p gpurun_out
timeout 900 python tools/bench_cut.py --k 21 --frames 50 > gpurun_out/cut_k21.json 2> gpurun_out/cut_k21.err; echo "cut21 exit $?"; tail -3 gpurun_out/cut_k21.err
timeout 1200 python tools/bench_cut.py --k 24 --frames 20 > gpurun_out/cut_k24.json 2> gpurun_out/cut_k24.err; echo "cut24 exit $?"; tail -3 gpurun_out/cut_k24.err
cat gpurun_out/cut_k21.json; echo; cat gpurun_out/cut_k24.json
