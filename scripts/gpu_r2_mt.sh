#!/bin/bash
# NOTE: the switch / kernel variant this script measures is not in the library any more: apply profiles/r02_mt2_experiment.patch to re-run it
# k_gemm_tc with two weight-row tiles per CTA (MT = 2) and SM-count-sized K splits: parity, then config 4 and the prefill A/B (QTC_GEMM_MT=1 forces one tile)
TAG=${1:-mt}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_batch.py tests/test_gpu_fused.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
timeout 900 python -m pytest tests/test_gpu_fullshape_parity.py -x -q > $OUT/pytest_full.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_full.log
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json; tail -3 $OUT/bench_batch64.err
QTC_GEMM_MT=1 timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64_mt1.json 2> $OUT/bench_batch64_mt1.err; echo "batch exit $?"; cat $OUT/bench_batch64_mt1.json; tail -3 $OUT/bench_batch64_mt1.err
timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_b1.json 2> $OUT/bench_b1.err; echo "bench exit $?"; cat $OUT/bench_b1.json; tail -3 $OUT/bench_b1.err
