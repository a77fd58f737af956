#!/bin/bash
# the contract's N > 1 launch on 2 GPUs of one box: both arms of the default line (config 2 replicas + config 5 split as `batched`)
TAG=${1:-multi2}; N=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > $OUT/ref_n$N.json 2> $OUT/ref_n$N.err; echo "ref N=$N exit $?"; cat $OUT/ref_n$N.json; tail -2 $OUT/ref_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench N=$N exit $?"; cat $OUT/bench_n$N.json; tail -3 $OUT/bench_n$N.err
