set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/pytest_s14.txt; cat gpurun_out/pytest_s14.txt
timeout 600 python tools/bench_eval_path.py --k 24 --frames 1000 > gpurun_out/eval_path_s14_n1.json 2> gpurun_out/eval_path_s14_n1.err; echo "eval exit $?"; tail -n 5 gpurun_out/eval_path_s14_n1.err; cat gpurun_out/eval_path_s14_n1.json
timeout 300 python tools/bench_dp_train.py > gpurun_out/dp_train_s14_ours_n1.json 2> gpurun_out/dp_train_s14_ours_n1.err; echo "dp ours exit $?"; tail -n 5 gpurun_out/dp_train_s14_ours_n1.err; cat gpurun_out/dp_train_s14_ours_n1.json
timeout 300 python tools/bench_dp_train.py --reference > gpurun_out/dp_train_s14_reference_n1.json 2> gpurun_out/dp_train_s14_reference_n1.err; echo "dp ref exit $?"; tail -n 5 gpurun_out/dp_train_s14_reference_n1.err; cat gpurun_out/dp_train_s14_reference_n1.json
timeout 600 python tools/bench_eval_path.py --k 25 --frames 200 --check-every 20 > gpurun_out/eval_path_s14_k25_n1.json 2> gpurun_out/eval_path_s14_k25_n1.err; echo "eval k25 exit $?"; tail -n 5 gpurun_out/eval_path_s14_k25_n1.err; cat gpurun_out/eval_path_s14_k25_n1.json
timeout 600 python bench.py > gpurun_out/bench_s14_ours.json 2> gpurun_out/bench_s14_ours.err; echo "bench exit $?"; tail -n 3 gpurun_out/bench_s14_ours.err; cut -c1-1500 gpurun_out/bench_s14_ours.json
