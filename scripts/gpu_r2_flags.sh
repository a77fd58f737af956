#!/bin/bash
# NOTE: flag 256 became the only path and QTC_PREFILL_STAGES was dropped after this run (profiles/README.md)
# A/B on one box: cluster kernel with the down projection's partials added straight into the fixed-point accumulator (QTC_FUSED_FLAGS=256:
# no cluster barrier / DSMEM fold in phase B), and the prefill GEMM with 2 pipeline stages (QTC_PREFILL_STAGES=2)
TAG=${1:-flags}; OUT=gpurun_out/$TAG; mkdir -p $OUT
QTC_FUSED_FLAGS=256 timeout 900 python -m pytest tests/test_gpu_cluster.py tests/test_gpu_fullshape_parity.py tests/test_gpu_fused.py -x -q > $OUT/pytest_f256.log 2>&1; echo "pytest f256 exit $?"; tail -3 $OUT/pytest_f256.log
QTC_PREFILL_STAGES=2 timeout 900 python -m pytest tests/test_gpu_fused.py -x -q > $OUT/pytest_pst2.log 2>&1; echo "pytest pst2 exit $?"; tail -3 $OUT/pytest_pst2.log
for v in "0 5" "256 5" "0 2" "256 2"; do set -- $v
  QTC_FUSED_FLAGS=$1 QTC_PREFILL_STAGES=$2 timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_f$1_p$2.json 2> $OUT/bench_f$1_p$2.err; echo "bench flags=$1 prefill_stages=$2 exit $?"; cat $OUT/bench_f$1_p$2.json
done
