#!/bin/bash
TAG=${1:-kv}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q --durations=10 > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -25 $OUT/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
