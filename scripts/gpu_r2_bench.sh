#!/bin/bash
# the bench lines the driver runs (N=1 default, reference arm) + the per-config lines
TAG=${1:-bench}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "default exit $?"; cat $OUT/bench_default.json; tail -3 $OUT/bench_default.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"; cat $OUT/bench_ref.json; tail -3 $OUT/bench_ref.err
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json; tail -3 $OUT/bench_batch64.err
timeout 600 python bench.py --model 0.6b --frames 74 --greedy --steps 3 --warmup 3 --no-cpu > $OUT/bench_06b.json 2> $OUT/bench_06b.err; echo "0.6b exit $?"; cat $OUT/bench_06b.json
