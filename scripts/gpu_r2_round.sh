#!/bin/bash
# checkpoint of the tree: whole GPU suite, then the batched line with the GEMM's pipeline depth A/B (QTC_GEMM_STAGES=2: two CTAs per SM)
TAG=${1:-round}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_gpu.log
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json; tail -3 $OUT/bench_batch64.err
QTC_GEMM_STAGES=2 timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64_st2.json 2> $OUT/bench_batch64_st2.err; echo "batch exit $?"; cat $OUT/bench_batch64_st2.json; tail -3 $OUT/bench_batch64_st2.err
QTC_GEMM_STAGES=2 timeout 600 python -m pytest tests/test_gpu_batch.py -x -q > $OUT/pytest_st2.log 2>&1; echo "pytest st2 exit $?"; tail -3 $OUT/pytest_st2.log
