#!/bin/bash
# skinny-GEMM prefill (R <= 64 prompt rows through qtc_gemm_tc_parts): parity tests, then the bench's latency object with the path on / off
TAG=${1:-prefill}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py tests/test_gpu_cluster.py tests/test_gpu_fullshape_parity.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_skinny.json 2> $OUT/bench_skinny.err; echo "bench exit $?"; cat $OUT/bench_skinny.json; tail -3 $OUT/bench_skinny.err
QTC_PREFILL_SKINNY=0 timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_conv.json 2> $OUT/bench_conv.err; echo "bench exit $?"; cat $OUT/bench_conv.json; tail -3 $OUT/bench_conv.err
