#!/bin/bash
# checkpoint of the tree: whole GPU suite, the driver's default bench line (config 2 + config 4 as `batched`), phase clocks of the cluster kernel
TAG=${1:-round}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "default exit $?"; cat $OUT/bench_default.json; tail -3 $OUT/bench_default.err
timeout 300 python scripts/phase_profile_cluster.py 1.7b 190 3 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
timeout 300 python scripts/phase_profile_cluster.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
