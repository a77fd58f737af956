#!/bin/bash
# streaming codec + backend flag: tests, then the default bench line (latency object)
TAG=${1:-st}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_stream.py tests/test_benchmark_cuda.py tests/test_cli_switch.py -x -q --durations=8 > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -18 $OUT/pytest.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fused.py tests/test_gpu_fullsize.py -x -q > $OUT/pytest2.log 2>&1; echo "pytest2 exit $?"; tail -5 $OUT/pytest2.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
