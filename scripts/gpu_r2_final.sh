#!/bin/bash
# closing run of the round: whole GPU suite, the driver's two arms, the per-config lines
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"; cat $OUT/bench_ref.json; tail -2 $OUT/bench_ref.err
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "default exit $?"; cat $OUT/bench_default.json; tail -2 $OUT/bench_default.err
timeout 600 python bench.py --model 0.6b --frames 74 --greedy --steps 3 --warmup 3 --no-cpu > $OUT/bench_06b.json 2> $OUT/bench_06b.err; echo "0.6b exit $?"; cat $OUT/bench_06b.json
timeout 600 python bench.py --codec-frames 3600 --steps 3 --warmup 3 --no-cpu > $OUT/bench_codec3600.json 2> $OUT/bench_codec3600.err; echo "codec exit $?"; cat $OUT/bench_codec3600.json
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json
timeout 900 python bench.py --queue 256 --frames 250 --steps 1 --warmup 1 --no-cpu > $OUT/bench_queue256.json 2> $OUT/bench_queue256.err; echo "queue exit $?"; cat $OUT/bench_queue256.json
