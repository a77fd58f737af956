#!/bin/bash
# cluster kernel: every CTA of a cluster walks the whole context (<= 512 positions) instead of an eighth + record exchange; QTC_FUSED_FLAGS=512 (at the time of the run: 512 = the split path; the switch was inverted afterwards, the split path is the default)
TAG=${1:-fullatt}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_cluster.py tests/test_gpu_fused.py tests/test_gpu_fullshape_parity.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
QTC_FUSED_FLAGS=512 timeout 900 python -m pytest tests/test_gpu_fused.py -x -q -k long_context > $OUT/pytest_split.log 2>&1; echo "pytest split exit $?"; tail -3 $OUT/pytest_split.log
for v in 0 512; do
  QTC_FUSED_FLAGS=$v timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_f$v.json 2> $OUT/bench_f$v.err; echo "bench flags=$v exit $?"
  python - $OUT/bench_f$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4), 'p99', d['latency']['frame_ms_p99'])
P
done
