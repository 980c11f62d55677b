set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 tools/bench_eval_path.py --k 24 --frames 1000 > gpurun_out/eval_path_s15_n$N.json 2> gpurun_out/eval_path_s15_n$N.err; echo "eval exit $?"; cat gpurun_out/eval_path_s15_n$N.json
timeout 300 $TR --master-port 29532 tools/bench_dp_train.py > gpurun_out/dp_train_s15_ours_n$N.json 2> gpurun_out/dp_train_s15_ours_n$N.err; echo "dp ours exit $?"; tail -n 3 gpurun_out/dp_train_s15_ours_n$N.err; cat gpurun_out/dp_train_s15_ours_n$N.json
timeout 300 $TR --master-port 29533 tools/bench_dp_train.py --no-overlap > gpurun_out/dp_train_s15_ours_noov_n$N.json 2> gpurun_out/dp_train_s15_ours_noov_n$N.err; echo "dp noov exit $?"; cat gpurun_out/dp_train_s15_ours_noov_n$N.json
