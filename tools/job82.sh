set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 tools/bench_eval_path.py --k 24 --frames 1000 > gpurun_out/eval_path_s17_n$N.json 2> gpurun_out/eval_path_s17_n$N.err; echo "eval exit $?"; cat gpurun_out/eval_path_s17_n$N.json
