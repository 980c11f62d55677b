set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for v in flat-bucket coalesce; do
timeout 300 $TR --master-port 29533 tools/bench_dp_train.py --no-overlap --$v > gpurun_out/dp_train_s16_${v}_n$N.json 2> gpurun_out/dp_train_s16_${v}_n$N.err; echo "dp $v exit $?"; tail -n 2 gpurun_out/dp_train_s16_${v}_n$N.err; cat gpurun_out/dp_train_s16_${v}_n$N.json
done
