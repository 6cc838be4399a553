set -x
O=gpurun_out
mkdir -p $O
N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py > $O/r1_dp_check_g$N.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu > $O/r1_bench_tf32_g$N.json 2>$O/bench_g$N.err
tail -12 $O/r1_dp_check_g$N.log; cat $O/r1_bench_tf32_g$N.json | cut -c1-400
