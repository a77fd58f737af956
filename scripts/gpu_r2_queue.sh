#!/bin/bash
# continuous batching (generate_queue), back-to-back batch codec decode, sub-talker input by table in the batched path: tests + config 4 A/B
TAG=${1:-queue}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_batch.py tests/test_gpu_paged_kv.py -x -q > $OUT/pytest_batch.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_batch.log
timeout 900 python -m pytest tests/test_gpu_fullshape_parity.py -x -q -k "batch" > $OUT/pytest_full_batch.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_full_batch.log
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json; tail -3 $OUT/bench_batch64.err
QTC_BATCH_PTAB=0 timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64_noptab.json 2> $OUT/bench_batch64_noptab.err; echo "batch exit $?"; cat $OUT/bench_batch64_noptab.json; tail -3 $OUT/bench_batch64_noptab.err
timeout 300 python bench.py --codec-frames 250 --steps 5 --warmup 3 --no-cpu > $OUT/bench_codec250.json 2> $OUT/bench_codec250.err; echo "codec exit $?"; cat $OUT/bench_codec250.json
