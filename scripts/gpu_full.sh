#!/bin/bash
# full-size property tests + config 1 line.  usage: gpurun --timeout 1500 -- 'bash scripts/gpu_full.sh <tag>'
TAG=${1:-full}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_fullsize.py -x -q > $OUT/pytest_full.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest_full.log
timeout 600 python bench.py --model 0.6b --frames 74 --greedy --steps 3 --warmup 3 > $OUT/bench_cfg1.json 2> $OUT/bench_cfg1.err; echo "cfg1 exit $?"; cat $OUT/bench_cfg1.json; tail -3 $OUT/bench_cfg1.err
