#!/bin/bash
# closing run of the round: whole GPU suite, smoke(), the driver's two arms
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"; cat $OUT/bench_ref.json; tail -2 $OUT/bench_ref.err
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "default exit $?"; cat $OUT/bench_default.json; tail -2 $OUT/bench_default.err
