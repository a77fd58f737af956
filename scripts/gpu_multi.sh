#!/bin/bash
# the contract's N > 1 launch line on N GPUs of one box: default (batch-1 replicas) and the lock-step batch (config 5 shape).
# usage: gpurun --gpus 2 --timeout 1200 -- 'bash scripts/gpu_multi.sh <tag> 2'
TAG=${1:-multi}; N=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench N=$N exit $?"; cat $OUT/bench_n$N.json; tail -3 $OUT/bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --batch 64 --frames 100 --steps 2 --warmup 1 --no-cpu > $OUT/batch_n$N.json 2> $OUT/batch_n$N.err; echo "batch N=$N exit $?"; cat $OUT/batch_n$N.json; tail -3 $OUT/batch_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > $OUT/ref_n$N.json 2> $OUT/ref_n$N.err; echo "ref N=$N exit $?"; cat $OUT/ref_n$N.json; tail -3 $OUT/ref_n$N.err
timeout 300 python bench.py --model 0.6b --frames 74 --greedy --steps 3 --warmup 3 > $OUT/bench_cfg1.json 2> $OUT/bench_cfg1.err; echo "cfg1 exit $?"; cat $OUT/bench_cfg1.json; tail -3 $OUT/bench_cfg1.err
